"""Generates tests/golden/iuwt_decomposition.npz from the reference's OWN
iuwt_decomposition.{h,cc} (oracle/_ref, built from /root/reference). Run in the dev container."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
from test_iuwt_oracle import image

assert oracle.ref_lib() is not None
img = image(96, 112, 2024)
n_scales = 3
coef = oracle.ref_iuwt_decompose(img, n_scales, True)
masks = (np.random.default_rng(9).random((n_scales, 96, 112)) > 0.5).astype(np.uint8)
rec = oracle.ref_iuwt_recompose(coef, n_scales, False, masks)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "iuwt_decomposition.npz"), image=img,
                    n_scales=n_scales, coefficients=coef, masks=masks, recomposed_masked=rec)
print("written")
