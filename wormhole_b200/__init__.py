"""wormhole_b200: the sparse minibatch training hot path of dmlc wormhole
(learn/linear, shared by learn/difacto) as hand-written sm_100a CUDA kernels behind a
C-ABI (include/wormhole_b200.h).  No CPU fallback: everything that computes needs the
built extension wormhole_b200/libwormhole_b200.so and a GPU."""
from .rowblock import RowBlock, LocalBlock, read_libsvm  # noqa: F401
