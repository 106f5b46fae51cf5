"""configs[0] inputs: ALL 39 of the reference's bundled test images (test_images/*.jpg) decoded the way the reference
does (PIL ``convert('RGB')`` + ``resize((256, 256))``, extract_rg_embeddings.py:260-261) and stored as uint8 arrays in
tests/golden/test_images_256.npz (build container only: /root/reference and PIL's JPEG decoder are used here; the GPU
box reads the committed arrays).  The float image the reference derives from these bytes (ToTensor / Normalize /
de-normalise / clip, :263-276) is recomputed by region_graph.denormalised_image in the tests; this script also checks
that helper against torchvision itself on every image."""
import glob
import os
import sys

import numpy as np
from PIL import Image

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
SRC = "/root/reference/test_images"
paths = sorted(glob.glob(os.path.join(SRC, "*.jpg")))
arrs = {os.path.basename(p).split(".")[0]: np.asarray(Image.open(p).convert("RGB").resize((256, 256)), dtype=np.uint8) for p in paths}

from torchvision import transforms  # noqa: E402

from camouflage_multimodal_b200.region_graph import denormalised_image  # noqa: E402

tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
for p in paths:
    t = tf(Image.open(p).convert("RGB").resize((256, 256))).numpy().transpose(1, 2, 0)
    ref = np.clip(t * np.array([0.229, 0.224, 0.225]) + np.array([0.485, 0.456, 0.406]), 0, 1)
    assert np.array_equal(ref, denormalised_image(arrs[os.path.basename(p).split(".")[0]])), p
out = os.path.join(ROOT, "tests", "golden", "test_images_256.npz")
np.savez_compressed(out, **arrs)
print("wrote", out, len(arrs), "images", os.path.getsize(out), "bytes; denormalised_image == torchvision on all")
