"""Golden vectors transcribed from the reference's own tests (data only).

Source: /root/reference/tensorflow_compression/python/layers/entropy_models_test.py
  G1 EntropyBottleneck   :179-199  (layer: channels_first, filters=(), init_scale=2 -> offset -2)
  G2 GaussianConditional :509-533  (scale_table 2**-5..2**0)
  G3 LogisticConditional :544-574
  G4 LaplacianConditional:585-615
All at precision=16, overflow_width=4 (entropy_models.py:235-239).
"""
import numpy as np

# ---- G1 ------------------------------------------------------------------------------
EB_BITSTRINGS = [
    b"\x91\xf4\xdan2\xd3q\x97\xd0\x91N1~\xc4\xb0;\xd38\xa8\x90",
    b"?\xc7\xf9\x17\xa8\xcfu\x99\x1e4\xfe\xe0\xd3U`z\x15v",
]
EB_CDF = np.array([
    [0, 5170, 11858, 19679, 27812, 35302, 65536],
    [0, 6100, 13546, 21671, 29523, 36269, 65536],
    [0, 6444, 14120, 22270, 29929, 36346, 65536],
], dtype=np.int32)
EB_CDF_LENGTH = np.array([7, 7, 7], dtype=np.int32)
EB_OFFSET = np.array([-2, -2, -2], dtype=np.int32)  # quantiles init (-2, 0, 2): minima = 2
EB_EXPECTED = np.array([
    [[-3., 2., 1., -3., -1., -3., -4., -2., 2.],
     [-2., 2., 4., 1., 0., -3., -3., 2., 4.],
     [1., 2., 4., -1., -3., 4., 0., -2., -3.]],
    [[0., 4., 0., 2., 4., 1., -2., 1., 4.],
     [2., 2., 3., -3., 4., -1., -1., 0., -1.],
     [3., 0., 3., -3., 3., 3., -3., -4., -1.]],
], dtype=np.float32)
# channels_first, per batch element (C=3, L=9): index = channel id
EB_INDEX = np.repeat(np.arange(3, dtype=np.int32)[:, None], 9, axis=1)

# ---- G2..G4 --------------------------------------------------------------------------
COND_INDEXES = np.array([
    [1, 2, 3, 4, 2, 2, 1],
    [5, 5, 1, 5, 3, 2, 3],
], dtype=np.int32)
COND_EXPECTED = np.array([
    [-3., 2., 1., -3., -1., -3., -4.],
    [-2., 2., -2., 2., 4., 1., 0.],
], dtype=np.float32)
COND_SCALE_TABLE = [2.0 ** x for x in range(-5, 1)]


def _pad(rows):
  w = max(len(r) for r in rows)
  return np.array([r + [0] * (w - len(r)) for r in rows], dtype=np.int32)


GAUSSIAN = dict(
    bitstrings=[
        b"\xff\xff\x13\xff\xff\x0f\xff\xef\xa9\x000\xb9\xffT\x87\xffUB",
        b"\x10\xf1m-\xf0r\xac\x97\xb6\xd5",
    ],
    cdf=_pad([
        [0, 1, 65534, 65535, 65536],
        [0, 1, 65534, 65535, 65536],
        [0, 2, 65533, 65535, 65536],
        [0, 1491, 64044, 65535, 65536],
        [0, 88, 10397, 55138, 65447, 65535, 65536],
        [0, 392, 4363, 20205, 45301, 61143, 65114, 65506, 65536],
    ]),
    cdf_length=np.array([5, 5, 5, 5, 7, 9], dtype=np.int32),
)
LOGISTIC = dict(
    bitstrings=[
        b"\xff\xff\x13\xff\xff\x0e\x17\xfd\xb5B\x03\xff\xf4\x11",
        b",yh\x13)\x12F\xfb",
    ],
    cdf=_pad([
        [0, 1, 65534, 65535, 65536],
        [0, 22, 65513, 65535, 65536],
        [0, 1178, 64357, 65535, 65536],
        [0, 159, 7809, 57721, 65371, 65530, 65536],
        [0, 52, 431, 3100, 17617, 47903, 62420, 65089, 65468, 65520, 65536],
        [0, 62, 230, 683, 1884, 4935, 11919, 24706, 40758, 53545, 60529, 63580, 64781,
         65234, 65402, 65464, 65536],
    ]),
    cdf_length=np.array([5, 5, 5, 7, 11, 17], dtype=np.int32),
)
LAPLACIAN = dict(
    bitstrings=[
        b"\xff\xff\x13\xff\xff\x0e\xea\xc1\xd9n'\xff\xfe*",
        b"\x1b\x9c\xd3\x06\xde_\xc0$",
    ],
    cdf=_pad([
        [0, 1, 65534, 65535, 65536],
        [0, 11, 65524, 65535, 65536],
        [0, 600, 64935, 65535, 65536],
        [0, 80, 4433, 61100, 65453, 65533, 65536],
        [0, 191, 1602, 12025, 53451, 63874, 65285, 65476, 65536],
        [0, 85, 315, 940, 2640, 7262, 19825, 45612, 58175, 62797, 64497, 65122, 65352, 65437,
         65536],
    ]),
    cdf_length=np.array([5, 5, 5, 7, 9, 15], dtype=np.int32),
)
for _g in (GAUSSIAN, LOGISTIC, LAPLACIAN):
  # offset = -pmf_center, pmf_center = (cdf_length - 3) / 2  (entropy_models.py:827-829,883)
  _g["offset"] = -((_g["cdf_length"] - 3) // 2)
CONDITIONALS = {"gaussian": GAUSSIAN, "logistic": LOGISTIC, "laplacian": LAPLACIAN}
