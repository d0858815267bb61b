"""Mirror of `tensorflow_compression/python/ops/spectral_ops.py:31-74`."""
import numpy as np
import scipy.fft

__all__ = ["irdft_matrix"]


def _rfft_packed(x, axis):
  """Real FFT in the packed layout of scipy.fftpack.rfft: [y0, Re y1, Im y1, Re y2, ...]."""
  n = x.shape[axis]
  spec = np.moveaxis(scipy.fft.rfft(x, axis=axis), axis, -1)
  out = np.empty(spec.shape[:-1] + (n,), np.float64)
  out[..., 0] = spec[..., 0].real
  pairs = (n - 1) // 2
  out[..., 1:2 * pairs + 1:2] = spec[..., 1:pairs + 1].real
  out[..., 2:2 * pairs + 2:2] = spec[..., 1:pairs + 1].imag
  if n % 2 == 0:
    out[..., -1] = spec[..., n // 2].real
  return np.moveaxis(out, -1, axis)


def irdft_matrix(shape, dtype=np.float32):
  """Orthonormal matrix M (prod(shape) x prod(shape)) such that `M @ rdft` is the kernel whose
  real DFT coefficients (packed, sqrt(2)-weighted so that M is orthonormal) are `rdft`."""
  shape = tuple(int(s) for s in shape)
  size = int(np.prod(shape))
  m = np.identity(size, dtype=np.float64).reshape((size,) + shape)
  for axis, n in enumerate(shape):
    m = _rfft_packed(m, axis + 1)
    sl = [slice(None)] * (len(shape) + 1)
    sl[axis + 1] = slice(1, None) if n % 2 == 1 else slice(1, -1)
    m[tuple(sl)] *= np.sqrt(2)
  m /= np.sqrt(size)
  return m.reshape(size, size).astype(dtype)
