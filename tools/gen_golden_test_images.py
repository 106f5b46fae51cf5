"""configs[0] inputs: two of the reference's bundled test images (test_images/img13.jpg, img28.jpg) decoded the way the
reference does (PIL ``convert('RGB')`` + ``resize((256, 256))``, extract_rg_embeddings.py:260-261) and stored as uint8
arrays in tests/golden/test_images_256.npz (build container only: /root/reference and PIL's JPEG decoder are used here;
the GPU box reads the committed arrays)."""
import os

import numpy as np
from PIL import Image

SRC = "/root/reference/test_images"
names = ["img13.jpg", "img28.jpg"]
arrs = {n.split(".")[0]: np.asarray(Image.open(os.path.join(SRC, n)).convert("RGB").resize((256, 256)), dtype=np.uint8) for n in names}
out = os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "test_images_256.npz")
np.savez_compressed(out, **arrs)
print("wrote", out, {k: (v.shape, int(v.mean())) for k, v in arrs.items()}, os.path.getsize(out))
