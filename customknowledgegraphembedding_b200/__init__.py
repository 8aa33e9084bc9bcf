"""B200-native drop-in for the data-parallel hot path of NguyenThaiHoc1/CustomKnowledgeGraphEmbedding:
``KGEModel.forward`` scoring, the negative-sampling ``train_step`` and the filtered ``test_step`` ranking,
as hand-written sm_100a CUDA kernels behind a C-ABI (``include/kge_b200.h``).  See DESIGN.md."""
from .model import KGEModel, MODEL_NAMES  # noqa: F401
from .optim import FusedAdam  # noqa: F401

__all__ = ["KGEModel", "FusedAdam", "MODEL_NAMES"]
