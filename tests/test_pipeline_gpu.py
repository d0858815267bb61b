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
