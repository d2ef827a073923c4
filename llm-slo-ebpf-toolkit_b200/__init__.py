"""B200-native batch correlation + attribution path of llm-slo-ebpf-toolkit.

Host-side mirror of the reference's Go interfaces for this path (pkg/correlation,
pkg/otel/processor/ebpfcorrelator, pkg/attribution) over the C ABI of libllmslo_gpu.so
(include/llmslo_gpu.h).  All arithmetic of the path runs in hand-written CUDA kernels for sm_100a;
there is no CPU fallback: importing works anywhere, but every compute call needs the built library
and a CUDA device and raises otherwise.
"""
from . import _abi  # noqa: F401
from ._lib import Context, LsgError, default_context, lib_path, load_library  # noqa: F401

__all__ = ["Context", "LsgError", "default_context", "lib_path", "load_library"]
