"""poporon-b200: B200-native batched Reed-Solomon / BCH codec behind libpoporon's C99 API.

This Python package is a thin ctypes mirror of the C API (``include/poporon.h``,
``include/poporon/batch.h``, ``include/poporon_cuda.h``) used by the tests and the benchmark.  All
work is done by two native libraries built in-tree under ``libpoporon_b200/lib``:

* ``libpoporon_cuda.so`` -- hand-written sm_100a kernels + the ``extern "C"`` device ABI
* ``libpoporon.so``      -- the C99 host library exporting the ``poporon_*`` API

There is no Python or CPU implementation of the codecs here: importing :mod:`libpoporon_b200.api`
fails loudly when the native libraries are missing, and handle creation fails without a GPU.
"""
from .build import build_native, lib_dir  # noqa: F401

__all__ = ["build_native", "lib_dir"]
