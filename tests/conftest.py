import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)


def pytest_configure(config):
  config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def pytest_collection_modifyitems(config, items):
  import torch
  if torch.cuda.is_available():
    return
  skip = pytest.mark.skip(reason="no CUDA device")
  for item in items:
    if "gpu" in item.keywords:
      item.add_marker(skip)


@pytest.fixture(scope="session", params=["reference", "port"])
def oracle_any(request):
  """Each CPU checker in turn; 'reference' is skipped where it was never built."""
  from oracle import oracle_lib
  if request.param == "reference" and not (
      os.path.exists(oracle_lib.REF_SO) or os.path.isdir("/root/reference")):
    pytest.skip("oracle/_ref not built and /root/reference absent")
  return oracle_lib.load(request.param)


@pytest.fixture(scope="session")
def oracle():
  """The strongest checker available (reference build if present, else the port)."""
  from oracle import oracle_lib
  return oracle_lib.load("best")


@pytest.fixture(scope="session")
def tfc():
  """The product package (needs the built libtfc_b200.so)."""
  import compression_b200 as tfc_mod
  return tfc_mod
