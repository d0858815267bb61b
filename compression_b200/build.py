"""Builds libtfc_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libtfc_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O3", "--expt-relaxed-constexpr",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
] + os.environ.get("TFC_NVCC_EXTRA", "").split()   # e.g. -DTFC_A0_KNOCKOUT for tools/conv_bench.py --debug


def _stale(obj, srcs):
  if not os.path.exists(obj):
    return True
  t = os.path.getmtime(obj)
  return any(os.path.getmtime(s) > t for s in srcs)


def _flags_changed(obj):
  """An object remembers the nvcc flags it was built with (obj + ".flags"): a build with other
  flags (TFC_NVCC_EXTRA) rebuilds it instead of silently linking the old variant."""
  try:
    return open(obj + ".flags").read() != " ".join(FLAGS)
  except OSError:
    return True


def build(verbose=False, force=False):
  srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
  hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
  objdir = os.path.join(HERE, "build")
  os.makedirs(objdir, exist_ok=True)
  objs, procs = [], []
  for s in srcs:
    o = os.path.join(objdir, os.path.basename(s)[:-3] + ".o")
    objs.append(o)
    if force or _stale(o, [s] + hdrs) or _flags_changed(o):
      cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
      procs.append((s, o, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
  failed = False
  for s, o, p in procs:
    out, _ = p.communicate()
    if p.returncode != 0 or verbose:
      sys.stderr.write(out.decode())
    if p.returncode != 0:
      failed = True
    else:
      with open(o + ".flags", "w") as f:
        f.write(" ".join(FLAGS))
  if failed:
    raise RuntimeError("nvcc failed")
  if force or procs or _stale(OUT, objs):
    subprocess.check_call([NVCC, "-shared", "-o", OUT] + objs + ["-lcudart_static", "-lcuda"]
                          if False else
                          [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
                           "-o", OUT] + objs + ["-lcuda"])
  return OUT


if __name__ == "__main__":
  print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
