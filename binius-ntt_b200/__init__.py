"""binius-ntt_b200 -- B200 (sm_100a) additive NTT over binary tower fields.

Host-side Python mirror of the reference's `ulvt` NTT surface (AdditiveNTTConf / AdditiveNTT /
NTTData, ntt/nttconf.cuh:9-60, ntt/additive_ntt.cuh:176-270) over the C ABI of libb2ntt.so
(include/b2ntt.h).  torch is used for device memory and streams only.  There is no CPU fallback:
importing works anywhere, computing needs the CUDA library and a GPU and fails loudly otherwise.
"""
from .api import (  # noqa: F401
    B2NTTError,
    AdditiveNTT,
    AdditiveNTTConf,
    BabyBearNTT,
    DataOrder,
    NTTData,
    GeneralNTT,
    MultiGPU,
    Plan,
    Sumcheck,
    bb31_decode,
    bb31_encode,
    bitslice_,
    check_gpu_capabilities,
    lib,
    lib_path,
    mul,
    mul_repeat,
)
from .multi import ColumnShardedRS, FourStepNTT, shard_columns  # noqa: F401
