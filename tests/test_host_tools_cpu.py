"""Host-side tooling that measurements depend on: the build script must not link an object built
with other nvcc flags (that is how a tracer build once passed for a plain one), and the
DRAM-traffic file bench.py reads for `roofline.traffic` is made from an ncu raw CSV."""
import csv
import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(path, name):
  spec = importlib.util.spec_from_file_location(name, path)
  mod = importlib.util.module_from_spec(spec)
  spec.loader.exec_module(mod)
  return mod


def test_build_objects_remember_their_flags(tmp_path):
  build = _load(os.path.join(ROOT, "compression_b200", "build.py"), "tfc_build_script")
  obj = str(tmp_path / "x.o")
  open(obj, "w").close()
  assert build._flags_changed(obj)                       # no record: rebuild
  with open(obj + ".flags", "w") as f:
    f.write(" ".join(build.FLAGS))
  assert not build._flags_changed(obj)
  with open(obj + ".flags", "w") as f:
    f.write(" ".join(build.FLAGS + ["-DTFC_A0_STAMPS=0"]))
  assert build._flags_changed(obj)                       # built as another variant: rebuild


def test_conv_traffic_from_ncu_raw_csv(tmp_path):
  tool = os.path.join(ROOT, "tools", "conv_traffic.py")
  mod = _load(tool, "tfc_conv_traffic")
  tensor = mod.TENSOR
  hdr = ["ID", "Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum",
         "dram__bytes_write.sum", tensor]
  units = ["", "", "ms", "Gbyte", "Mbyte", "%"]
  rows = []
  for i, name in enumerate(mod.LAYERS):
    for launch in range(2):
      # the first launch of a layer (cold, weight repack) must be ignored
      rows.append([str(2 * i + launch), "void <unnamed>::k%d(int)" % i,
                   "9.0" if launch == 0 else "0.5", "1.0", "250", "50"])
  path = tmp_path / "raw.csv"
  with open(path, "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(hdr)
    w.writerow(units)
    w.writerows(rows)
  out = subprocess.run([sys.executable, tool, str(path), "test build"], capture_output=True,
                       text=True, check=True).stdout
  d = json.loads(out)
  n_hs = sum(1 for name in mod.LAYERS if name.startswith("HS"))
  launches = len(mod.LAYERS) + n_hs                      # hyper-synthesis runs twice per step
  assert launches == 17
  assert abs(d["dram_bytes_per_step"] - launches * 1.25e9) < 1.0
  assert abs(d["ncu_ms_per_step_serialised"] - launches * 0.5) < 1e-9
  assert d["per_layer"]["A0"]["kernel"] == "k0" and d["build"] == "test build"


def test_bench_reads_the_newest_traffic_file():
  """bench.py takes roofline.traffic from the newest profiles/r*_conv_traffic.json: that file must
  exist, name its command, and hold the per-layer numbers it sums."""
  files = sorted(f for f in os.listdir(os.path.join(ROOT, "profiles")) if f.endswith("_conv_traffic.json"))
  assert files, "no conv traffic file under profiles/"
  d = json.load(open(os.path.join(ROOT, "profiles", files[-1])))
  assert "ncu --set full" in d["command"]
  total = sum((2 if k.startswith("HS") else 1) * v["dram_bytes"] for k, v in d["per_layer"].items())
  assert abs(total - d["dram_bytes_per_step"]) <= 1e-6 * total
