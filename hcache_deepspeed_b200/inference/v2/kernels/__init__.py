from .core_ops import (BlasLibLinear, CUDABiasActivation, CUDAFPLN, CUDAFPPostLN, CUDAFPPreLN, CUDAGatedActivation,  # noqa: F401
                       CUDARMSNorm, CUDARMSPreNorm)
from .cutlass_ops import MixedGEMM, MoEGEMM  # noqa: F401
from .ds_kernel import DSKernelBase  # noqa: F401
from .ragged_ops import (AtomBuilder, BlockedFlashAttn, BlockedRotaryEmbeddings, LinearBlockedKVCopy, MoEGather, MoEScatter,  # noqa: F401
                         RaggedEmbeddingKernel, RaggedLogitsGather, RaggedTopKGating)
