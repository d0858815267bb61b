"""TEST INFRASTRUCTURE ONLY — CPU restatement (torch-CPU / numpy, float32 or float64) of the
float half of the hot path.  Never imported by the product package.

Parity status: the reference's Python needs TensorFlow 1.15, which cannot run in this image,
so this restatement is pinned only by the properties and golden tables the reference's own
tests hold (tests/test_oracle_layers.py): scipy-convolve equivalence and identity-kernel
alignment (signal_conv_test.py:32-183), GDN closed forms (gdn_test.py:49-75), the golden CDF
tables of entropy_models_test.py:519-526,554-567,595-608, likelihood-vs-scipy
(entropy_models_test.py:448-463).  Float parity with TF's own kernels beyond those
tolerances is UNPINNED.

Each function cites the reference lines it follows (relative to
/root/reference/tensorflow_compression/python/).
"""
import numpy as np
import scipy.special
import scipy.stats
import torch
import torch.nn.functional as F


# ---- ops/padding_ops.py:27-56 --------------------------------------------------------------
def same_padding_for_kernel(shape, corr, strides_up=None):
  rank = len(shape)
  if strides_up is None:
    strides_up = rank * (1,)
  if corr:
    padding = [(s // 2, (s - 1) // 2) for s in shape]
  else:
    padding = [((s - 1) // 2, s // 2) for s in shape]
  return [((padding[i][0] - 1) // strides_up[i] + 1, (padding[i][1] - 1) // strides_up[i] + 1)
          for i in range(rank)]


# ---- layers/signal_conv.py:709-814 (2-D, same_zeros|valid, channels_last) --------------------
def signal_conv2d(x, kernel, bias=None, corr=False, strides_down=1, strides_up=1,
                  padding="same_zeros", extra_pad_end=True, dtype=torch.float64):
  """x [N,H,W,Cin], kernel [kh,kw,Cin,Cout] (numpy or torch) -> [N,Ho,Wo,Cout] torch `dtype`.

  Follows the reference literally: the odd-kernel flips (:723-738), then either
  tf.nn.conv2d with explicit padding (:487-518) or tf.nn.conv2d_backprop_input with explicit
  padding followed by [::strides_down] slicing (:629-707)."""
  x = torch.as_tensor(np.asarray(x)).to(dtype).permute(0, 3, 1, 2)        # NCHW
  k = torch.as_tensor(np.asarray(kernel)).to(dtype)
  kh, kw = k.shape[:2]
  su, sd = int(strides_up), int(strides_down)
  odd = kh % 2 == 1 and kw % 2 == 1
  if not corr and su == 1 and odd:
    corr, k = True, torch.flip(k, (0, 1))
  elif corr and su != 1 and odd:
    corr, k = False, torch.flip(k, (0, 1))
  prepad = [(0, 0), (0, 0)]
  if padding == "valid":
    pad = [(0, 0), (0, 0)]
  else:
    pad = same_padding_for_kernel((kh, kw), corr, (su, su))
    if padding == "same_reflect":
      # signal_conv.py:752-758: pre-pad with tf.pad(REFLECT), then the VALID ops
      x = F.pad(x, (pad[1][0], pad[1][1], pad[0][0], pad[0][1]), mode="reflect")
      prepad, pad = pad, [(0, 0), (0, 0)]
  if corr and su == 1:
    xp = F.pad(x, (pad[1][0], pad[1][1], pad[0][0], pad[0][1]))
    out = F.conv2d(xp, k.permute(3, 2, 0, 1), stride=sd)                  # cross-correlation
  elif not corr:
    # conv2d_backprop_input(out_shape, kernel^T, x, strides=su, explicit padding p):
    # gradient of a stride-su correlation == full transposed conv cropped by p at the start
    if padding == "valid":
      p = [(kh - 1, kh - 1), (kw - 1, kw - 1)]
    else:       # signal_conv.py:657-663, pre-padding folded into the explicit padding
      p = [(prepad[0][0] * su + kh // 2, prepad[0][1] * su + (kh - 1) // 2),
           (prepad[1][0] * su + kw // 2, prepad[1][1] * su + (kw - 1) // 2)]
    n, _, h, w = x.shape
    def length(l, s, kk, ps):
      return l * s + ((kk - 1) - ps) if extra_pad_end else l * s + ((kk - 1) - (s - 1) - ps)
    ho, wo = length(h, su, kh, sum(p[0])), length(w, su, kw, sum(p[1]))
    # F.conv_transpose2d weight layout [Cin, Cout, kh, kw]; it computes
    # full[t] = sum_{i,a: su*i + a = t} x[i] * w[a]  (size (l-1)*su + k)
    full = F.conv_transpose2d(x, k.permute(2, 3, 0, 1), stride=su)
    full = F.pad(full, (0, max(0, p[1][0] + wo - full.shape[3]),
                        0, max(0, p[0][0] + ho - full.shape[2])))
    out = full[:, :, p[0][0]:p[0][0] + ho, p[1][0]:p[1][0] + wo]
    if sd > 1:
      out = out[:, :, ::sd, ::sd]
  else:
    raise NotImplementedError("corr with upsampling and an even kernel (signal_conv.py:393)")
  out = out.permute(0, 2, 3, 1)
  if bias is not None:
    out = out + torch.as_tensor(np.asarray(bias)).to(dtype)
  return out.contiguous()


# ---- layers/gdn.py:171-206 ------------------------------------------------------------------
def gdn(x, beta, gamma, inverse=False, dtype=torch.float64):
  x = torch.as_tensor(np.asarray(x)).to(dtype)
  beta = torch.as_tensor(np.asarray(beta)).to(dtype)
  gamma = torch.as_tensor(np.asarray(gamma)).to(dtype)
  norm = torch.matmul(x * x, gamma) + beta          # gamma indexed [input k, output c]
  return x * (torch.sqrt(norm) if inverse else torch.rsqrt(norm))


# ---- layers/parameterizers.py:175-191 --------------------------------------------------------
def nonnegative_reparam(stored, minimum=0.0, reparam_offset=2 ** -18):
  pedestal = reparam_offset ** 2
  bound = (minimum + pedestal) ** .5
  return np.maximum(np.asarray(stored, np.float64), bound) ** 2 - pedestal


# ---- layers/entropy_models.py:402-436 ----------------------------------------------------------
def eb_logits_cumulative(x, matrices, biases, factors):
  """x: [C, 1, B] float64; matrices[i]: softplus'ed [C, f_{i+1}, f_i]; factors tanh'ed."""
  logits = x
  for i in range(len(matrices)):
    logits = np.matmul(matrices[i], logits) + biases[i]
    if i < len(factors):
      logits = logits + factors[i] * np.tanh(logits)
  return logits


def _sigmoid(x):
  return scipy.special.expit(x)


def eb_params64(matrices_raw, biases, factors_raw):
  """Stored variables -> reparameterised float64 (softplus / tanh, :474, :488)."""
  m = [np.logaddexp(0, np.asarray(a, np.float64)) for a in matrices_raw]
  b = [np.asarray(a, np.float64) for a in biases]
  f = [np.tanh(np.asarray(a, np.float64)) for a in factors_raw]
  return m, b, f


def eb_call(x, medians, matrices, biases, factors, likelihood_bound=1e-9):
  """EntropyBottleneck.call(training=False) (:148-187, :609-614, :626-658); x [..., C]."""
  x32 = np.asarray(x, np.float32)
  med = np.asarray(medians, np.float32)
  values = np.floor(x32 + (np.float32(.5) - med)) + med           # float32, as the reference
  c = x32.shape[-1]
  v = values.astype(np.float64).reshape(-1, c).T[:, None, :]      # (C, 1, batch)
  lower = eb_logits_cumulative(v - .5, matrices, biases, factors)
  upper = eb_logits_cumulative(v + .5, matrices, biases, factors)
  sign = -np.sign(lower + upper)
  lik = np.abs(_sigmoid(sign * upper) - _sigmoid(sign * lower))
  lik = lik[:, 0, :].T.reshape(x32.shape)
  return values, np.maximum(lik, likelihood_bound)


def eb_symbols(x, medians):
  x32 = np.asarray(x, np.float32)
  med = np.asarray(medians, np.float32)
  return np.floor(x32 + (np.float32(.5) - med)).astype(np.int32)


def eb_tables(quantiles, matrices, biases, factors, pmf_to_cdf, precision=16):
  """EntropyBottleneck table construction (:521-559, :588-596) in float32 inputs / float64
  density evaluation.  pmf_to_cdf: callable(pmf_row float32 [n]) -> int32 [n+1]."""
  q = np.asarray(quantiles, np.float32)
  medians = q[:, 0, 1]
  minima = np.maximum(np.ceil(medians - q[:, 0, 0]).astype(np.int32), 0)
  maxima = np.maximum(np.ceil(q[:, 0, 2] - medians).astype(np.int32), 0)
  offset = -minima
  pmf_start = medians - minima.astype(np.float32)
  pmf_length = maxima + minima + 1
  max_length = int(pmf_length.max())
  samples = np.arange(max_length, dtype=np.float32)[None, None, :] + pmf_start[:, None, None]
  s64 = samples.astype(np.float64)
  lower = eb_logits_cumulative(s64 - .5, matrices, biases, factors)
  upper = eb_logits_cumulative(s64 + .5, matrices, biases, factors)
  sign = -np.sign(lower + upper)
  pmf = np.abs(_sigmoid(sign * upper) - _sigmoid(sign * lower))[:, 0, :].astype(np.float32)
  tail = (_sigmoid(lower[:, 0, :1]) + _sigmoid(-upper[:, 0, -1:])).astype(np.float32)
  cdf = np.zeros((q.shape[0], max_length + 2), np.int32)
  for r in range(q.shape[0]):
    n = int(pmf_length[r])
    cdf[r, :n + 2] = pmf_to_cdf(np.concatenate([pmf[r, :n], tail[r]]).astype(np.float32))
  return cdf, (pmf_length + 2).astype(np.int32), offset.astype(np.int32), medians


# ---- layers/entropy_models.py:800-885, :965-972 --------------------------------------------------
def _cumulative(kind, t):
  t = np.asarray(t)
  if kind == "gaussian":
    return (t.dtype.type(.5) * scipy.special.erfc(t.dtype.type(-(2 ** -0.5)) * t)).astype(t.dtype)
  if kind == "logistic":
    return (t.dtype.type(1) / (t.dtype.type(1) + np.exp(-t))).astype(t.dtype)
  if kind == "laplacian":
    e = np.exp(-np.abs(t))
    return (np.where(t > 0, 2 - e, e) / 2).astype(t.dtype)
  raise ValueError(kind)


_PPF = {"gaussian": scipy.stats.norm.ppf, "logistic": scipy.stats.logistic.ppf,
        "laplacian": scipy.stats.laplace.ppf}


def conditional_tables(scale_table, pmf_to_cdf, kind="gaussian", tail_mass=2 ** -8):
  """SymmetricConditional.build table part (:827-865, :883), float32 like the TF graph."""
  table = np.array(sorted(float(s) for s in scale_table), np.float64)
  multiplier = -_PPF[kind](tail_mass / 2)
  center = np.ceil(table * multiplier).astype(int)
  length = 2 * center + 1
  max_length = int(length.max())
  samples = np.abs(np.arange(max_length, dtype=int) - center[:, None]).astype(np.float32)
  sc = table.astype(np.float32)[:, None]
  upper = _cumulative(kind, ((np.float32(.5) - samples) / sc).astype(np.float32))
  lower = _cumulative(kind, ((np.float32(-.5) - samples) / sc).astype(np.float32))
  pmf = (upper - lower).astype(np.float32)
  tail = (np.float32(2) * lower[:, :1]).astype(np.float32)
  cdf = np.zeros((len(table), max_length + 2), np.int32)
  for r in range(len(table)):
    n = int(length[r])
    cdf[r, :n + 2] = pmf_to_cdf(np.concatenate([pmf[r, :n], tail[r]]).astype(np.float32))
  return cdf, (length + 2).astype(np.int32), (-center).astype(np.int32)


def conditional_indexes(scale, scale_table, scale_bound=None):
  """:822-825 + :869-881 — literally the foldr: start at levels-1, subtract [scale <= s]."""
  table = np.array(sorted(float(s) for s in scale_table), np.float32)
  s = np.asarray(scale, np.float32)
  bound = table[0] if scale_bound is None else np.float32(scale_bound)
  if bound > 0:
    s = np.maximum(s, bound)
  idx = np.full(s.shape, len(table) - 1, np.int32)
  for t in table[:-1]:
    idx -= (s <= t).astype(np.int32)
  return idx


def conditional_call(y, scale, scale_table, kind="gaussian", scale_bound=None,
                     likelihood_bound=1e-9):
  """SymmetricConditional.call(training=False), mean=None (:148-187, :895-903, :916-932)."""
  table0 = np.float32(sorted(scale_table)[0])
  values = np.floor(np.asarray(y, np.float32) + np.float32(.5))
  s = np.asarray(scale, np.float32)
  bound = table0 if scale_bound is None else np.float32(scale_bound)
  if bound > 0:
    s = np.maximum(s, bound)
  v = np.abs(values).astype(np.float64)
  s = s.astype(np.float64)
  lik = _cumulative(kind, (.5 - v) / s) - _cumulative(kind, (-.5 - v) / s)
  return values, np.maximum(lik, likelihood_bound)


def conditional_symbols(y):
  return np.floor(np.asarray(y, np.float32) + np.float32(.5)).astype(np.int32)
