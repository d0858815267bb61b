"""End-to-end parity of the bmshj2018 pipeline on the GPU against the CPU restatement with the
SAME weights and tables (SURVEY §8d "parity checks")."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup(oracle):
  from compression_b200.examples import bmshj2018
  from oracle import cpu_codec
  model = bmshj2018.Bmshj2018Model(192, "cuda", seed=0)
  cpu = cpu_codec.CpuBmshj2018(model.export_weights(), oracle, threads=8,
                               tables=model.export_tables())
  g = torch.Generator().manual_seed(3)
  x = torch.rand((2, 80, 112, 3), generator=g, dtype=torch.float32)   # 80, 112: ragged /16
  return model, cpu, x


def _psnr(a, b):
  return 10 * np.log10(1.0 / np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))


def test_roundtrip_and_bit_exact_strings(setup, oracle):
  model, cpu, x = setup
  xd = x.cuda()
  c = model.compress(xd)
  assert c.y_shape == (5, 7) and c.z_shape == (2, 2)        # ceil division on the way down
  host = model.to_host(c)
  y, z, sigma = model.analyse(xd)
  gc = model._conditional(sigma)
  y_sym, y_idx, _ = gc._symbols_and_indexes(y.contiguous())
  t = model.export_tables()
  # (1) same symbols + indexes + tables -> same bytes as the reference coder
  for b in range(2):
    want = oracle.ui_encode(y_sym[b].cpu().numpy(), y_idx[b].cpu().numpy(), t["gc_cdf"],
                            t["gc_len"], t["gc_off"], 16, 4)
    assert host.string[b] == want
  z_sym = model.entropy_bottleneck._quantize_symbols(z).cpu().numpy()
  z_idx = np.broadcast_to(np.arange(192, dtype=np.int32), z_sym.shape[1:])
  for b in range(2):
    assert host.side_string[b] == oracle.ui_encode(z_sym[b], z_idx, t["eb_cdf"], t["eb_len"],
                                                   t["eb_off"], 16, 4)
  # (2) the decoder, which re-derives sigma from the side string, recovers the same symbols
  x_hat = model.decompress(model.to_device(host))
  nf = 192
  z_hat = model.entropy_bottleneck.decompress_device(c.side_string, c.z_shape + (nf,), nf,
                                                     torch.bfloat16)
  sigma2 = model.hyper_synthesis_transform(z_hat)[:, :5, :7, :].contiguous()
  assert torch.equal(sigma2, sigma), "decoder-side sigma must equal encoder-side bit for bit"
  y_hat = model._conditional(sigma2).decompress_device(c.string)
  assert torch.equal(y_hat.to(torch.int32), y_sym)
  assert tuple(x_hat.shape) == (2, 80, 112, 3)


def test_cpu_decoder_reads_gpu_strings_and_psnr_delta(setup):
  """The CPU pipeline decodes the GPU's strings (interchange), and with the SAME symbols the
  two synthesis paths agree to |dPSNR| <= 0.01 dB (north-star tolerance)."""
  model, cpu, x = setup
  c = model.compress(x.cuda())
  host = model.to_host(c)
  from compression_b200.python.ops import range_coding_ops as rc

  def slab(strings):
    cap = max(len(s) for s in strings) + 8
    out = np.zeros((len(strings), cap), np.uint8)
    for i, s in enumerate(strings):
      out[i, :len(s)] = np.frombuffer(s, np.uint8)
    return out, np.array([len(s) for s in strings], np.int64)

  cc = dict(string=slab(host.string), side_string=slab(host.side_string), x_shape=c.x_shape,
            y_shape=c.y_shape, z_shape=c.z_shape)
  # z symbols decode identically; the CPU's float32 sigma may index a few rows differently
  # from the GPU's bf16-operand sigma, so y is decoded with the GPU's indexes here
  y, z, sigma = model.analyse(x.cuda())
  gc = model._conditional(sigma)
  y_sym, y_idx, _ = gc._symbols_and_indexes(y.contiguous())
  t = model.export_tables()
  n = y_idx[0].numel()
  dec = cpu.coder.ui_decode_batch(cc["string"][0], cc["string"][1],
                                  y_idx.cpu().numpy().reshape(2, -1), n, t["gc_cdf"],
                                  t["gc_len"], t["gc_off"], threads=2)
  np.testing.assert_array_equal(dec.reshape(y_sym.shape), y_sym.cpu().numpy())
  # same symbols -> both synthesis transforms
  y_hat = y_sym.float()
  x_gpu = model.synthesis_transform(y_hat.to(torch.bfloat16))[:, :80, :112].float().cpu().numpy()
  x_cpu = cpu._run(__import__("oracle.cpu_codec", fromlist=["SYNTHESIS"]).SYNTHESIS,
                   y_hat.cpu().numpy())[:, :80, :112].numpy()
  d = abs(_psnr(x.numpy(), x_gpu) - _psnr(x.numpy(), x_cpu))
  assert d <= 0.01, d
  rel = np.abs(x_gpu - x_cpu).max() / (np.abs(x_cpu).max() + 1e-9)
  assert rel < 0.05, rel


def test_analysis_close_to_cpu(setup):
  """Transforms (bf16 activations between layers) against the float32 CPU restatement."""
  model, cpu, x = setup
  y, z, sigma = model.analyse(x.cuda())
  y_c, z_sym_c, sigma_c = cpu.analyse(x.numpy())
  scale = np.abs(y_c).max()
  assert np.abs(y.cpu().numpy() - y_c).max() <= 0.03 * scale
  agree = np.mean(np.floor(y.cpu().numpy() + .5) == np.floor(y_c + .5))
  assert agree > 0.97, agree


def test_results_independent_of_batch_position(setup):
  """H4: an image codes to the same bytes whatever batch it is in."""
  model, cpu, x = setup
  a = model.to_host(model.compress(x.cuda()))
  b = model.to_host(model.compress(x[1:].cuda()))
  assert a.string[1] == b.string[0] and a.side_string[1] == b.side_string[0]


def test_sharded_batches_code_to_the_same_strings(setup):
  """the batch is split contiguously across ranks with no collective; whatever the
  number of shards, every image must code to the bytes of the unsharded run and decode to the
  same pixels (here the shards run one after the other on one GPU)."""
  from compression_b200 import parallel
  model, cpu, _ = setup
  g = torch.Generator().manual_seed(9)
  x = torch.rand((5, 64, 96, 3), generator=g, dtype=torch.float32).cuda()
  full = model.to_host(model.compress(x))
  x_hat_full = model.decompress(model.to_device(full))
  for world in (2, 3, 5):
    strings, side, recon = [], [], []
    for lo, hi in parallel.shard_bounds(5, world):
      part = model.to_host(model.compress(x[lo:hi]))
      strings += part.string
      side += part.side_string
      recon.append(model.decompress(model.to_device(part)))
    assert strings == full.string and side == full.side_string
    assert torch.equal(torch.cat(recon), x_hat_full)


@pytest.mark.parametrize("depth", [1, 3])
def test_pipeline_matches_batch_by_batch(setup, depth):
  """The multi-stream pipeline (packed coder CTAs, chunked transform kernels, deferred status
  checks) returns exactly what compress()/decompress() return batch by batch."""
  from compression_b200.examples import bmshj2018
  model, cpu, _ = setup
  g = torch.Generator().manual_seed(11)
  batches = [torch.rand((3, 48, 80, 3), generator=g, dtype=torch.float32) for _ in range(5)]
  want = []
  for x in batches:
    c = model.compress(x.cuda())
    want.append((model.to_host(c), model.decompress(c)))
  with model.pipeline(depth=depth) as pipe:
    got = pipe.roundtrip_device([x.cuda() for x in batches])
    for (host, x_hat), g_hat in zip(want, got):
      assert torch.equal(x_hat, g_hat)
    u8 = [(x * 255).round().to(torch.uint8).pin_memory() for x in batches]
    seen = []
    # the callback's strings and pixels are views of pinned staging buffers: copy them out
    pipe.roundtrip_host(u8, lambda k, c, xh: seen.append(
        (k, list(c.string), list(c.side_string), c.string.total_bytes, xh.clone())))
    listed = pipe.roundtrip_host(u8[:2])
  assert [k for k, _, _, _, _ in seen] == list(range(5))
  for k, string, side_string, total, xh in seen:
    ref = model.compress(u8[k].cuda().float() / 255)
    ref_host = model.to_host(ref)
    assert string == ref_host.string and side_string == ref_host.side_string
    assert total == sum(len(b) for b in ref_host.string)
    if k < 2:
      assert listed[k][0].string == string and torch.equal(listed[k][1], xh)
    assert torch.equal(xh, bmshj2018.quantize_image(model.decompress(ref)).cpu())


@pytest.mark.parametrize("back_lag,host_slots", [(1, None), (3, None), (2, 7), (6, 7), (9, 10)])
def test_pipeline_host_waits_do_not_change_results(setup, back_lag, host_slots):
  """roundtrip_host's two host waits (back_lag batches after a compress for its strings;
  host_slots - back_lag batches after a decompress for the slot's buffers) only move when the
  host blocks: strings and pixels are those of the batch-by-batch calls whatever the schedule,
  also when the lag exceeds the batches in flight and when slots are reused several times."""
  from compression_b200.examples import bmshj2018
  model, _, _ = setup
  g = torch.Generator().manual_seed(23)
  u8 = [torch.randint(0, 256, (2, 40, 56, 3), generator=g, dtype=torch.uint8).pin_memory()
        for _ in range(17)]
  with model.pipeline(depth=3, back_lag=back_lag, host_slots=host_slots) as pipe:
    assert pipe.back_lag < pipe.host_slots
    seen = []
    pipe.roundtrip_host(u8, lambda k, c, xh: seen.append(
        (k, list(c.string), list(c.side_string), xh.clone())))
  assert [k for k, _, _, _ in seen] == list(range(len(u8)))
  for k, string, side_string, xh in seen:
    ref = model.compress(u8[k].cuda().float() / 255)
    ref_host = model.to_host(ref)
    assert string == ref_host.string and side_string == ref_host.side_string
    assert torch.equal(xh, bmshj2018.quantize_image(model.decompress(ref)).cpu())


def test_tfci_files_round_trip_one_image_each(setup, tmp_path):
  """every image leaves as a `.tfci` file in the reference's container
  (bmshj2018.py:339-347) and decodes from it (`:369-377`) to what the batch decode produces."""
  model, _, x = setup
  xd = x.cuda()
  c = model.compress(xd)
  host = model.to_host(c)
  want = model.decompress(c)
  for b in range(x.shape[0]):
    path = tmp_path / ("img%d.tfci" % b)
    path.write_bytes(model.pack_tfci(host, image=b, model_id="bmshj2018-b200"))
    one = model.unpack_tfci(path.read_bytes())
    assert one.string == [host.string[b]] and one.side_string == [host.side_string[b]]
    assert (one.x_shape, one.y_shape, one.z_shape) == (host.x_shape, host.y_shape, host.z_shape)
    got = model.decompress(model.to_device(one))
    assert torch.equal(got[0], want[b])


def test_cluster_multicast_conv_gives_the_same_strings(setup):
  """TFC_CFG_CONV_CLUSTER: CTA pairs sharing the weight tiles by TMA multicast change nothing."""
  import compression_b200 as tfc
  model, _, x = setup
  xd = x.cuda()
  base = model.to_host(model.compress(xd))
  try:
    tfc.configure(conv_cluster=2)
    paired = model.to_host(model.compress(xd))
    x_hat = model.decompress(model.to_device(paired))
  finally:
    tfc.configure(conv_cluster=0)
  assert paired.string == base.string and paired.side_string == base.side_string
  assert torch.equal(x_hat, model.decompress(model.to_device(base)))


# ---- a model that really reconstructs (the random-init PSNR test was vacuous) -------
def _structured_model(seed=0, quant=24.0, jitter=8e-4):
  """bmshj2018 wiring with identity-preserving kernels: the analysis transform is three
  space-to-depth steps and a scaled 2x2 average, the synthesis transform its (pseudo-)inverse,
  GDN / IGDN are the identity (gamma = 0, beta = 1), and EVERY weight carries a small random
  jitter so that all taps and channels contribute.  PSNR of smooth images is ~27-30 dB."""
  from compression_b200.examples import bmshj2018
  from compression_b200.python.layers import gdn as gdn_lib
  model = bmshj2018.Bmshj2018Model(192, "cuda", seed=seed)
  g = torch.Generator().manual_seed(seed + 1)

  def put(layer, w):
    w = w + jitter * float(w.abs().max()) * torch.randn(w.shape, generator=g)
    kp = layer.kernel_parameterizer
    wd = w.to(layer.kernel_var.device)
    layer.kernel_var = kp.init(wd) if kp is not None else wd
    if isinstance(layer.activation, gdn_lib.GDN):
      act = layer.activation
      c = act.num_channels
      act.reparam_gamma = act.gamma_parameterizer.init(torch.zeros((c, c), device=wd.device))
      act.reparam_beta = act.beta_parameterizer.init(torch.ones((c,), device=wd.device))
    layer.invalidate()

  a, s = model.analysis_transform.layers, model.synthesis_transform.layers
  cins = [3, 12, 48]
  for li, cin in enumerate(cins):                    # space to depth: 3 -> 12 -> 48 -> 192
    w = torch.zeros((5, 5, a[li].input_channels, 192))
    for da in range(2):
      for db in range(2):
        for c in range(cin):
          w[2 + da, 2 + db, c, (da * 2 + db) * cin + c] = 1.0
    put(a[li], w)
  w = torch.zeros((5, 5, 192, 192))
  for da in range(2):
    for db in range(2):
      w[2 + da, 2 + db, torch.arange(192), torch.arange(192)] = quant / 4
  put(a[3], w)
  w = torch.zeros((5, 5, 192, 192))                  # nearest-neighbour upsampling / quant
  for di in range(2):
    for dj in range(2):
      w[2 + di, 2 + dj, torch.arange(192), torch.arange(192)] = 1.0 / quant
  put(s[0], w)
  for li, cout in ((1, 48), (2, 12), (3, 3)):        # depth to space: 192 -> 48 -> 12 -> 3
    w = torch.zeros((5, 5, 192, s[li].filters))
    for di in range(2):
      for dj in range(2):
        for o in range(cout):
          w[2 + di, 2 + dj, (di * 2 + dj) * cout + o, o] = 1.0
    put(s[li], w)
  return model


def _smooth_images(batch, h, w, seed):
  rng = np.random.default_rng(seed)
  yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
  out = np.zeros((batch, h, w, 3), np.float32)
  for b in range(batch):
    for c in range(3):
      f = rng.uniform(1 / 160.0, 1 / 90.0, size=2)
      ph = rng.uniform(0, 2 * np.pi, size=2)
      out[b, :, :, c] = 0.5 + 0.22 * np.sin(2 * np.pi * f[0] * yy + ph[0]) * \
          np.cos(2 * np.pi * f[1] * xx + ph[1]) + 0.1 * (xx / w - 0.5)
  return np.clip(out, 0, 1)


def test_reconstructing_model_psnr_and_per_transform_l2(oracle):
  """With weights that reconstruct (PSNR > 25 dB) the GPU transforms (bf16 operands and
  activations) must agree with the float32 CPU restatement to relative L2 <= 1e-2 per transform,
  and with the SAME symbols fed to both synthesis paths the PSNR may move by at most 0.01 dB
  (north-star tolerance) — a wrong tile tail, tap or channel shows up here as a PSNR drop."""
  from oracle import cpu_codec
  model = _structured_model()
  cpu = cpu_codec.CpuBmshj2018(model.export_weights(), oracle, threads=8,
                               tables=model.export_tables())
  x = _smooth_images(2, 144, 208, 5)                 # 144, 208: ragged tiles at every level
  xd = torch.from_numpy(x).cuda()

  def rel_l2(got, want):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    return float(np.sqrt(((got - want) ** 2).sum() / ((want ** 2).sum() + 1e-30)))

  y, z, sigma = model.analyse(xd)
  y_c, z_sym_c, sigma_c = cpu.analyse(x)
  assert rel_l2(y.cpu().numpy(), y_c) <= 1e-2
  # hyper-analysis on the SAME input (|y| of the CPU), hyper-synthesis on the SAME z symbols: a
  # z that rounds differently on the two sides would otherwise move sigma by a whole step
  z_gpu = model.hyper_analysis_transform(torch.from_numpy(np.abs(y_c)).cuda().to(torch.bfloat16))
  z_cpu = cpu._run(cpu_codec.HYPER_ANALYSIS, torch.abs(torch.from_numpy(y_c))).numpy()
  assert rel_l2(z_gpu.cpu().numpy(), z_cpu) <= 1e-2
  z_hat_c = z_sym_c.astype(np.float32) + cpu.medians
  s_gpu = model.hyper_synthesis_transform(torch.from_numpy(z_hat_c).cuda().to(torch.bfloat16))
  assert rel_l2(s_gpu[:, :y.shape[1], :y.shape[2]].cpu().numpy(), sigma_c) <= 1e-2
  assert np.mean(np.floor(y.cpu().numpy() + .5) == np.floor(y_c + .5)) > 0.9
  c = model.compress(xd)
  x_hat = model.decompress(c).cpu().numpy()
  psnr_e2e = _psnr(x, np.clip(x_hat, 0, 1))
  assert psnr_e2e > 25.0, psnr_e2e
  # same symbols into both synthesis transforms
  y_sym = torch.floor(y + .5)
  x_gpu = model.synthesis_transform(y_sym.to(torch.bfloat16))[:, :144, :208].float().cpu().numpy()
  x_cpu = cpu._run(cpu_codec.SYNTHESIS, y_sym.cpu().numpy())[:, :144, :208].numpy()
  assert rel_l2(x_gpu, x_cpu) <= 1e-2
  p_gpu, p_cpu = _psnr(x, x_gpu), _psnr(x, x_cpu)
  assert p_cpu > 25.0 and abs(p_gpu - p_cpu) <= 0.01, (p_gpu, p_cpu)
  # the uint8 output of the e2e leg is the quantised float reconstruction
  from compression_b200.examples import bmshj2018
  xq = model.decompress(c, out_dtype=torch.uint8)
  assert torch.equal(xq, bmshj2018.quantize_image(model.decompress(c)))
  # and the strings still are the CPU coder's strings of the GPU's symbols (escape-heavy here:
  # |y| reaches `quant` while the random hyperprior predicts small scales)
  host = model.to_host(c)
  gc = model._conditional(sigma)
  ys, yi, _ = gc._symbols_and_indexes(y.contiguous())
  t = model.export_tables()
  for b in range(2):
    assert host.string[b] == oracle.ui_encode(ys[b].cpu().numpy(), yi[b].cpu().numpy(),
                                              t["gc_cdf"], t["gc_len"], t["gc_off"], 16, 4)
