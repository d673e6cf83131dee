"""The four config values the corrupt -> denoise path and its callers read
(lib/model/config.py:63,66,72,160,169,172,252,255).  Not a port of the reference's
easydict `cfg`: only what `get_minibatch` / `_get_image_blob` / `test_net` use around the path."""
from types import SimpleNamespace

import numpy as np

cfg = SimpleNamespace(
    PIXEL_MEANS=np.array([[[102.9801, 115.9465, 122.7717]]]),     # config.py:252
    RNG_SEED=3,                                                     # config.py:255
    TRAIN=SimpleNamespace(SCALES=(600,), MAX_SIZE=1000, IMS_PER_BATCH=1, BATCH_SIZE=128, USE_ALL_GT=True),
    TEST=SimpleNamespace(SCALES=(600,), MAX_SIZE=1000, NMS=0.3),    # config.py:160,169,176
)
