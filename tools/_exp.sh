mkdir -p gpurun_out
L=gpurun_out/r2s_train_bf16.log
: > $L
timeout 900 python -m pytest tests/test_train_gpu.py -x -q -m gpu 2>&1 | tail -4 >> $L
timeout 300 python - >> $L 2>&1 <<'PY'
import torch, sys
sys.path.insert(0, '.')
from tests.test_train_gpu import _ref_losses
from compression_b200.examples import bls2017
model = bls2017.Bls2017Model(128, "cuda", seed=3)
trainer = bls2017.Bls2017Trainer(model, lmbda=0.01)
g = torch.Generator().manual_seed(4)
for shape, ys in (((4, 64, 64, 3), (4, 4, 4, 128)), ((8, 128, 128, 3), (8, 8, 8, 128))):
  x = torch.rand(shape, generator=g).cuda()
  noise = (torch.rand(ys, generator=g) - .5).cuda()
  trainer.main.zero_grad(); out = trainer.losses(x, noise); out["loss"].backward(); got = trainer.main.grad.clone()
  trainer.main.zero_grad(); ref = _ref_losses(trainer, x, noise); ref.backward(); want = trainer.main.grad.clone()
  cos = torch.dot(got, want) / (got.norm() * want.norm())
  print("shape", shape, "cosine", float(cos), "rel l2", float((got - want).norm() / want.norm()), "loss", float(out["loss"].detach()), float(ref.detach()))
PY
timeout 300 python tools/train_bench.py >> $L 2>&1
timeout 300 python tools/train_profile.py --top 14 2>&1 | tail -17 >> $L
cat $L
