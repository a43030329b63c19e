"""CNN layer kernels (tcgen05 implicit-GEMM conv, fused bias+ReLU+pool, dropout, linear).
Filled in by the CNN milestone; ``available()`` gates ``models.CNN`` onto them."""


def available():
    return False
