"""TEST INFRASTRUCTURE ONLY — the bmshj2018 hyperprior pipeline on the host CPU.

Restates /root/reference/examples/bmshj2018.py compress (:285-341) and decompress (:365-399)
with the float transforms of oracle/tfc_np.py (torch-CPU, float32) and the range coder of the
CPU checker (oracle/_ref = the reference's own C++ when built, else the port).  Used by tests
(parity of the GPU pipeline given the same weights) and by bench.py as the timed CPU baseline
and the `--impl reference` arm.  Labelled "CPU restatement, not TF": TensorFlow 1.15 cannot
run in this image (SURVEY.md probe table).
"""
import numpy as np
import torch

from oracle import tfc_np

SCALES_MIN, SCALES_MAX, SCALES_LEVELS = 0.11, 256, 64

# (name, kernel, corr, strides_down, strides_up, use_bias, activation) per bmshj2018.py:73-198
ANALYSIS = [("a%d" % i, 5, True, 2, 1, True, "gdn" if i < 3 else None) for i in range(4)]
SYNTHESIS = [("s%d" % i, 5, False, 1, 2, True, "igdn" if i < 3 else None) for i in range(4)]
HYPER_ANALYSIS = [("ha0", 3, True, 1, 1, True, "relu"), ("ha1", 5, True, 2, 1, True, "relu"),
                  ("ha2", 5, True, 2, 1, False, None)]
HYPER_SYNTHESIS = [("hs0", 5, False, 1, 2, True, "relu"), ("hs1", 5, False, 1, 2, True, "relu"),
                   ("hs2", 3, False, 1, 1, True, None)]


def random_weights(num_filters=192, seed=0):
  """Random-init weights with the reference's initialisers (variance_scaling fan_in truncated
  normal kernels, zero biases, beta=1, gamma=0.1*I, EB per entropy_models.py:461-515)."""
  rng = np.random.default_rng(seed)
  nf = num_filters
  w = {}

  def conv(name, k, cin, cout, bias, act):
    std = np.sqrt(1.0 / (k * k * cin)) / .87962566103423978
    kern = np.clip(rng.normal(size=(k, k, cin, cout)) * std, -2 * std, 2 * std)
    w[name + "/kernel"] = kern.astype(np.float32)
    if bias:
      w[name + "/bias"] = np.zeros(cout, np.float32)
    if act in ("gdn", "igdn"):
      w[name + "/beta"] = np.ones(cout, np.float32)
      w[name + "/gamma"] = (0.1 * np.eye(cout)).astype(np.float32)

  for spec, cin0, couts in ((ANALYSIS, 3, [nf] * 4), (SYNTHESIS, nf, [nf] * 3 + [3]),
                            (HYPER_ANALYSIS, nf, [nf] * 3), (HYPER_SYNTHESIS, nf, [nf] * 3)):
    cin = cin0
    for (name, k, _, _, _, bias, act), cout in zip(spec, couts):
      conv(name, k, cin, cout, bias, act)
      cin = cout
  f = (1, 3, 3, 3, 1)
  scale = 10 ** (1 / 4)
  for i in range(4):
    w["eb/matrix_%d" % i] = np.full((nf, f[i + 1], f[i]), np.log(np.expm1(1 / scale / f[i + 1])),
                                    np.float32)
    w["eb/bias_%d" % i] = rng.uniform(-.5, .5, size=(nf, f[i + 1], 1)).astype(np.float32)
    if i < 3:
      w["eb/factor_%d" % i] = np.zeros((nf, f[i + 1], 1), np.float32)
  w["eb/quantiles"] = np.tile(np.array([[[-10., 0., 10.]]], np.float32), (nf, 1, 1))
  return w


class CpuBmshj2018(object):

  def __init__(self, weights, coder, threads=1, tables=None):
    """coder: an oracle_lib.Oracle.  tables: optional dict with the integer tables to use
    (eb_cdf, eb_len, eb_off, gc_cdf, gc_len, gc_off) — parity tests pass the GPU model's so
    that both sides code against identical tables; otherwise they are built here."""
    self.w = weights
    self.coder = coder
    self.threads = int(threads)
    self.nf = weights["a0/kernel"].shape[-1]
    torch.set_num_threads(max(1, self.threads))
    self.scale_table = np.exp(np.linspace(np.log(SCALES_MIN), np.log(SCALES_MAX), SCALES_LEVELS))
    nl = len([k for k in weights if k.startswith("eb/matrix_")])
    self.eb = tfc_np.eb_params64([weights["eb/matrix_%d" % i] for i in range(nl)],
                                 [weights["eb/bias_%d" % i] for i in range(nl)],
                                 [weights["eb/factor_%d" % i] for i in range(nl - 1)])
    self.medians = weights["eb/quantiles"][:, 0, 1].astype(np.float32)
    if tables is None:
      p2c = lambda p: coder.pmf_to_cdf(p[None], 16)[0]
      eb_cdf, eb_len, eb_off, _ = tfc_np.eb_tables(weights["eb/quantiles"], *self.eb, p2c)
      gc_cdf, gc_len, gc_off = tfc_np.conditional_tables(self.scale_table, p2c)
      tables = dict(eb_cdf=eb_cdf, eb_len=eb_len, eb_off=eb_off, gc_cdf=gc_cdf, gc_len=gc_len,
                    gc_off=gc_off)
    self.t = tables

  def _run(self, spec, x):
    for name, k, corr, sd, su, bias, act in spec:
      x = tfc_np.signal_conv2d(x, self.w[name + "/kernel"],
                               self.w.get(name + "/bias") if bias else None, corr, sd, su,
                               "same_zeros", True, dtype=torch.float32)
      if act == "relu":
        x = torch.relu(x)
      elif act in ("gdn", "igdn"):
        x = tfc_np.gdn(x, self.w[name + "/beta"], self.w[name + "/gamma"], act == "igdn",
                       dtype=torch.float32)
    return x

  def analyse(self, x):
    y = self._run(ANALYSIS, x)
    z = self._run(HYPER_ANALYSIS, torch.abs(y))
    z_sym = tfc_np.eb_symbols(z.numpy(), self.medians)
    z_hat = z_sym.astype(np.float32) + self.medians
    sigma = self._run(HYPER_SYNTHESIS, z_hat)[:, :y.shape[1], :y.shape[2], :]
    return y.numpy(), z_sym, sigma.numpy()

  def compress(self, x):
    y, z_sym, sigma = self.analyse(x)
    t = self.t
    b = y.shape[0]
    z_idx = np.broadcast_to(np.arange(self.nf, dtype=np.int32), z_sym.shape[1:]).copy()
    side, side_n = self.coder.ui_encode_batch(z_sym.reshape(b, -1), z_idx.reshape(-1),
                                              t["eb_cdf"], t["eb_len"], t["eb_off"],
                                              threads=self.threads)
    y_sym = tfc_np.conditional_symbols(y)
    y_idx = tfc_np.conditional_indexes(sigma, self.scale_table)
    string, string_n = self.coder.ui_encode_batch(y_sym.reshape(b, -1), y_idx.reshape(b, -1),
                                                  t["gc_cdf"], t["gc_len"], t["gc_off"],
                                                  threads=self.threads)
    return dict(string=(string, string_n), side_string=(side, side_n), x_shape=x.shape[1:3],
                y_shape=y.shape[1:3], z_shape=z_sym.shape[1:3], y_sym=y_sym, y_idx=y_idx,
                z_sym=z_sym)

  def decompress(self, c):
    t = self.t
    side, side_n = c["side_string"]
    b = side.shape[0]
    zs = tuple(c["z_shape"]) + (self.nf,)
    z_idx = np.broadcast_to(np.arange(self.nf, dtype=np.int32), zs).copy().reshape(-1)
    z_sym = self.coder.ui_decode_batch(side, side_n, z_idx, z_idx.size, t["eb_cdf"], t["eb_len"],
                                       t["eb_off"], threads=self.threads).reshape((b,) + zs)
    z_hat = z_sym.astype(np.float32) + self.medians
    sigma = self._run(HYPER_SYNTHESIS, z_hat)[:, :c["y_shape"][0], :c["y_shape"][1], :].numpy()
    y_idx = tfc_np.conditional_indexes(sigma, self.scale_table)
    string, string_n = c["string"]
    n = y_idx[0].size
    y_sym = self.coder.ui_decode_batch(string, string_n, y_idx.reshape(b, -1), n, t["gc_cdf"],
                                       t["gc_len"], t["gc_off"], threads=self.threads)
    y_hat = y_sym.reshape(y_idx.shape).astype(np.float32)
    x_hat = self._run(SYNTHESIS, y_hat)
    return x_hat[:, :c["x_shape"][0], :c["x_shape"][1], :].numpy(), y_sym.reshape(y_idx.shape)


# ---- bls2017 (BASELINE.json configs[0]: 128 filters, GDN, 256x256 batch 8 on the CPU) ---------
# (name, kernel, corr, strides_down, strides_up, use_bias, activation) per bls2017.py:70-127
BLS_ANALYSIS = [("a0", 9, True, 4, 1, True, "gdn"), ("a1", 5, True, 2, 1, True, "gdn"),
                ("a2", 5, True, 2, 1, False, None)]
BLS_SYNTHESIS = [("s0", 5, False, 1, 2, True, "igdn"), ("s1", 5, False, 1, 2, True, "igdn"),
                 ("s2", 9, False, 1, 4, True, None)]


def random_weights_bls2017(num_filters=128, seed=0):
  """Random-init weights of the bls2017 model with the reference's initialisers (see
  random_weights)."""
  rng = np.random.default_rng(seed)
  nf = num_filters
  w = {}
  cin = 3
  for spec, couts in ((BLS_ANALYSIS, [nf] * 3), (BLS_SYNTHESIS, [nf, nf, 3])):
    for (name, k, _, _, _, bias, act), cout in zip(spec, couts):
      std = np.sqrt(1.0 / (k * k * cin)) / .87962566103423978
      kern = np.clip(rng.normal(size=(k, k, cin, cout)) * std, -2 * std, 2 * std)
      w[name + "/kernel"] = kern.astype(np.float32)
      if bias:
        w[name + "/bias"] = np.zeros(cout, np.float32)
      if act in ("gdn", "igdn"):
        w[name + "/beta"] = np.ones(cout, np.float32)
        w[name + "/gamma"] = (0.1 * np.eye(cout)).astype(np.float32)
      cin = cout
  f = (1, 3, 3, 3, 1)
  scale = 10 ** (1 / 4)
  for i in range(4):
    w["eb/matrix_%d" % i] = np.full((nf, f[i + 1], f[i]), np.log(np.expm1(1 / scale / f[i + 1])),
                                    np.float32)
    w["eb/bias_%d" % i] = rng.uniform(-.5, .5, size=(nf, f[i + 1], 1)).astype(np.float32)
    if i < 3:
      w["eb/factor_%d" % i] = np.zeros((nf, f[i + 1], 1), np.float32)
  w["eb/quantiles"] = np.tile(np.array([[[-10., 0., 10.]]], np.float32), (nf, 1, 1))
  return w


class CpuBls2017(CpuBmshj2018):
  """/root/reference/examples/bls2017.py compress (:205-271) and decompress (:274-305) on the
  host: analysis transform, EntropyBottleneck symbols, one range-coded string per image; the
  decoder dequantises and runs the synthesis transform."""

  def __init__(self, weights, coder, threads=1, tables=None):
    self.w = weights
    self.coder = coder
    self.threads = int(threads)
    self.nf = weights["a0/kernel"].shape[-1]
    torch.set_num_threads(max(1, self.threads))
    nl = len([k for k in weights if k.startswith("eb/matrix_")])
    self.eb = tfc_np.eb_params64([weights["eb/matrix_%d" % i] for i in range(nl)],
                                 [weights["eb/bias_%d" % i] for i in range(nl)],
                                 [weights["eb/factor_%d" % i] for i in range(nl - 1)])
    self.medians = weights["eb/quantiles"][:, 0, 1].astype(np.float32)
    if tables is None:
      p2c = lambda p: coder.pmf_to_cdf(p[None], 16)[0]
      eb_cdf, eb_len, eb_off, _ = tfc_np.eb_tables(weights["eb/quantiles"], *self.eb, p2c)
      tables = dict(eb_cdf=eb_cdf, eb_len=eb_len, eb_off=eb_off)
    self.t = tables

  def compress(self, x):
    y = self._run(BLS_ANALYSIS, x).numpy()
    y_sym = tfc_np.eb_symbols(y, self.medians)
    b = y.shape[0]
    idx = np.broadcast_to(np.arange(self.nf, dtype=np.int32), y_sym.shape[1:]).copy()
    t = self.t
    string, string_n = self.coder.ui_encode_batch(y_sym.reshape(b, -1), idx.reshape(-1),
                                                  t["eb_cdf"], t["eb_len"], t["eb_off"],
                                                  threads=self.threads)
    return dict(string=(string, string_n), x_shape=x.shape[1:3], y_shape=y.shape[1:3],
                y_sym=y_sym)

  def decompress(self, c):
    t = self.t
    string, string_n = c["string"]
    b = string.shape[0]
    ys = tuple(c["y_shape"]) + (self.nf,)
    idx = np.broadcast_to(np.arange(self.nf, dtype=np.int32), ys).copy().reshape(-1)
    y_sym = self.coder.ui_decode_batch(string, string_n, idx, idx.size, t["eb_cdf"], t["eb_len"],
                                       t["eb_off"], threads=self.threads).reshape((b,) + ys)
    y_hat = y_sym.astype(np.float32) + self.medians
    x_hat = self._run(BLS_SYNTHESIS, y_hat)
    return x_hat[:, :c["x_shape"][0], :c["x_shape"][1], :].numpy(), y_sym
