"""bench.py's supplementary-measurement helper must never raise or block: the headline JSON line is
printed after it returns."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location('pdm_bench_module', os.path.join(ROOT, 'bench.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_run_json_lines_collects_and_degrades():
    b = _bench()
    ok = b.run_json_lines([sys.executable, '-c', 'print("noise"); print(\'{"impl": "engine", "ms_per_step": 1.5}\')'], 60)
    assert ok == [{'impl': 'engine', 'ms_per_step': 1.5}]
    bad = b.run_json_lines([sys.executable, '-c', 'import sys; print("boom", file=sys.stderr); sys.exit(3)'], 60)
    assert len(bad) == 1 and 'exit code 3' in bad[0]['unavailable'] and 'boom' in bad[0]['unavailable']
    slow = b.run_json_lines([sys.executable, '-c', 'import time; time.sleep(30)'], 1)
    assert slow == [{'unavailable': 'timed out after 1 s'}]
    gone = b.run_json_lines(['/nonexistent/interpreter'], 5)
    assert 'unavailable' in gone[0]


def test_bench_cli_contract():
    b = _bench()
    assert b.PER_GPU_BATCH == 512 and 'images/sec' in b.METRIC
    # flags the driver passes must parse; --no-cnn exists
    import argparse  # noqa: F401
    argv, sys.argv = sys.argv, ['bench.py', '--gpus', '1', '--steps', '5', '--warmup', '3', '--no-cnn', '--impl', 'reference']
    try:
        called = {}
        b.run_reference = lambda opts: called.update(vars(opts))
        b.main()
    finally:
        sys.argv = argv
    assert called['gpus'] == 1 and called['steps'] == 5 and called['warmup'] == 3 and called['cnn'] is False


def test_child_process_geometry(monkeypatch):
    b = _bench()
    monkeypatch.setenv('MASTER_PORT', '29611')
    monkeypatch.setenv('TORCHELASTIC_USE_AGENT_STORE', 'True')
    monkeypatch.setenv('GROUP_RANK', '0')
    env = b.child_env(3, 3, 8, b.child_port(1))
    assert env['RANK'] == '3' and env['LOCAL_RANK'] == '3' and env['WORLD_SIZE'] == '8'
    assert env['MASTER_ADDR'] == '127.0.0.1' and 'GROUP_RANK' not in env
    assert not any(k.startswith('TORCHELASTIC_') for k in env)
    ports = {b.child_port(k) for k in (1, 2, 5)}
    assert len(ports) == 3 and all(20000 <= p < 40000 and p != 29611 for p in ports)
    # every rank derives the same port from the same parent rendezvous
    assert b.child_port(2) == b.child_port(2)
