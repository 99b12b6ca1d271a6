"""pim-compression_b200 -- B200-native backend for PIM-compression's block-parallel Snappy codec.

The product is native: ``csrc/`` (sm_100a CUDA kernels + the C-ABI shim declared in
``include/pim_snappy.h``) and ``cli/dpu_snappy.c``.  This Python package is only the thin
ctypes binding that tests and bench.py use to call that C ABI; it contains no codec logic and
no CPU fallback (loading fails loudly if ``libpim_snappy.so`` has not been built).

Import name: ``pim_compression_b200`` (the directory name carries a hyphen; the repo root has a
one-file import shim).
"""
from . import api  # noqa: F401
from .api import (  # noqa: F401
    LIB_PATH,
    SNAPPY_OK,
    SNAPPY_INVALID_INPUT,
    SNAPPY_BUFFER_TOO_SMALL,
    SYNTH_TEXT,
    SYNTH_RANDOM,
    SYNTH_MIX,
    SYNTH_BIN,
    HostBufferContext,
    ProgramRuntime,
    SnappyError,
    build,
    lib,
    compress,
    decompress,
    compress_device,
    compress_device_noalloc,
    compress_batch_device,
    decompress_batch_device,
    decompress_device,
    gather_shards,
    block_offsets,
    synth_lib,
    synth_host,
    synth_device,
    set_devices,
    set_chunk_bytes,
    set_extended_blocks,
    shard_range,
    device_count,
    launch_count,
    profile,
    profile_read,
    PinnedBuffer,
)
