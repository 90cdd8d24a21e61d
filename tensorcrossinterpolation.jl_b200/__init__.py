"""B200-native TensorCI2 hot path behind the reference's own API names.

Host logic (sweep strategy, index-set bookkeeping, convergence control) is Python mirroring the Julia
package; every arithmetic step calls the C ABI of libtci_b200.so (hand-written sm_100a CUDA).  The
package is importable without a GPU (so CPU-only tests can check symbols), but any computation
raises if the library or a CUDA device is missing — there is no CPU fallback.
"""
from ._lib import Handle, TCIError, default_handle, lib, SO_PATH          # noqa: F401
from .matrixlu import (rrLU, rrlu, rrlu_device, submatrixargmax, rrlu_last_timing, arrlu, rrlu_function, Rng, solve,   # noqa: F401
                       pushrandomsubset)
from .dense import Mat, gemm, luci_factors, solve_right, trsm_right, svd, diag_scale                              # noqa: F401
from .matrixluci import MatrixLUCI                                                  # noqa: F401
from .batcheval import (BatchEvaluator, BuiltinFunction, LorentzFunction, SumFunction, QuanticsExpSin,          # noqa: F401
                        QuanticsExpPlusSin, QuanticsExp2, DenseFunction, HashFunction, UserFunction, filltensor,
                        _batchevaluate_dispatch, to_numpy)
from .tensortrain import TensorTrain, AbstractTensorTrain, DeviceCore, tensortrain, compress, _factorize, add, subtract, norm2, norm, _addtttensor                       # noqa: F401
from .tensorci2 import (TensorCI2, crossinterpolate2, optimize, convergencecriterion, kronecker_left,      # noqa: F401
                        kronecker_right, DefaultGlobalPivotFinder, GlobalPivotSearchInput, forwardsweep,
                        HostFunctionAdapter, ThreadedBatchEvaluator, makebatchevaluatable, optfirstpivot, ENGINES)
from .tensor import DTensor, _contract                                                                  # noqa: F401
from .contraction import (Contraction, ElementwiseMap, contract, contract_naive, contract_zipup, contract_TCI)   # noqa: F401
from .parallel import Comm, ShardedBatchEvaluator, partition, bond_passes, MailboxPool, gather_sharded_U                               # noqa: F401
from .cachedfunction import CachedFunction                                                              # noqa: F401
from .integration import integrate, kronrod                                                            # noqa: F401
from .cachedtensortrain import TTCache, estimatetrueerror, _floatingzone                                # noqa: F401
