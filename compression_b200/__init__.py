"""compression_b200 — B200-native (sm_100a) hot path of tensorflow_compression.

Mirrors `tensorflow_compression/__init__.py:31-40`, which star-imports the layers and ops into
one namespace (`tfc.*`).
"""
import os as _os

# The multi-batch Pipeline (examples/bmshj2018.py) keeps 12-20 CUDA streams busy; with the
# driver's default of 8 hardware work queues several streams share a queue and wait on each
# other (measured at 1080p: 1332 -> 1599 MPix/s with 32 queues).  Only takes effect if set before
# the CUDA context exists, which is the case when this package is imported before the first CUDA
# call; an explicit setting in the environment wins.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from compression_b200._lib import InvalidArgumentError, CudaError, launch_count, configure
from compression_b200.python.ops.range_coding_ops import (
    range_encode, range_decode, unbounded_index_range_encode, unbounded_index_range_decode,
    pmf_to_quantized_cdf, deferred_checks)
from compression_b200.python.ops.math_ops import lower_bound, upper_bound
from compression_b200.python.ops.padding_ops import same_padding_for_kernel
from compression_b200.python.ops.spectral_ops import irdft_matrix
from compression_b200.python.layers.parameterizers import (
    Parameterizer, StaticParameterizer, RDFTParameterizer, NonnegativeParameterizer)
from compression_b200.python.layers.gdn import GDN
from compression_b200.python.layers.signal_conv import SignalConv1D, SignalConv2D
from compression_b200.python.layers.entropy_models import (
    EntropyModel, EntropyBottleneck, SymmetricConditional, GaussianConditional,
    LogisticConditional, LaplacianConditional)
from compression_b200.python.util.packed_tensors import PackedTensors
