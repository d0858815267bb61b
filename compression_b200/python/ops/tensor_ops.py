"""Small fused elementwise helpers of the codec pipelines (no counterpart module in the
reference; these are the `abs(y)` of examples/bmshj2018.py:304 and dtype casts)."""
import ctypes

import torch

from compression_b200 import _lib


def abs_to_bf16(x):
  """|x| as bfloat16 — the input of the hyper-analysis transform."""
  _lib.require_cuda(x, "x")
  x = x.float().contiguous()
  out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
  _lib.check(_lib.lib().tfc_abs_to_bf16(_lib.ptr(x), ctypes.c_int64(x.numel()), _lib.ptr(out),
                                        _lib.stream_ptr(x.device)))
  return out
