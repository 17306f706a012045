"""orfanage_b200 — B200-native (sm_100a) ORFanage hot path.

The product is native code: ``liborfx.so`` (hand-written CUDA kernels behind the C-ABI of
``include/orfx.h``), ``liborfx_host.so`` (the C++ host front-end: GTF/FASTA loading, template
preparation, locus bundling, batching — ``include/orfx_pack.h``) and the ``bin/orfanage`` CLI that
keeps the reference's command-line surface (orfanage.cpp:831-1125).  This Python package is only
the thin ctypes plumbing the benchmark and the parity tests use to reach those libraries; it holds
no algorithm and there is no CPU fallback: every computing call fails loudly without the CUDA
library or without a GPU.
"""
from ._capi import (  # noqa: F401
    OrfxError,
    Params,
    Batch,
    Results,
    Context,
    Workload,
    load_device_library,
    load_host_library,
    device_library_path,
    host_library_path,
    cli_path,
    MODES,
    STATUS_NOTES,
    encode_protein,
    shard_pairs,
)

__all__ = [
    "OrfxError", "Params", "Batch", "Results", "Context", "Workload", "load_device_library",
    "load_host_library", "device_library_path", "host_library_path", "cli_path", "MODES",
    "STATUS_NOTES", "encode_protein", "shard_pairs",
]
