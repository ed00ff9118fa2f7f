"""B200-native batched Viterbi decoder (update + chainback) behind the reference's per-codec API.

The product is ``libb200viterbi.so`` (CUDA kernels for sm_100a + the C-ABI of ``include/b200_viterbi.h``);
this package only builds/loads it and mirrors the reference's decoder interface on the host.
"""
from .decoder import (  # noqa: F401
    B200ViterbiError, BatchDecoder, WyConfig, CODEC_BY_NAME, STANDARD_POLYS, EXPORTED_SYMBOLS,
    KA9Q27, KA9Q29, KA9Q615, KA9Q224, SPIRAL27, SPIRAL29, SPIRAL47, SPIRAL49, SPIRAL615, WY_U8, WY_U16,
    load_library, library_path, symbol_dtype, metric_dtype, shard_range, shard_range_weighted, ShardedDecoder, decode_noalloc_dev, NOALLOC_SCRATCH_BYTES, StreamDecoder, PunctureSegment,
)
from .build import build_library  # noqa: F401
