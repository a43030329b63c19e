from .timing import DeviceTimer, max_over_ranks, nvtx_range

__all__ = ['DeviceTimer', 'max_over_ranks', 'nvtx_range']
