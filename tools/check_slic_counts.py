"""Known-answer check of the SLIC oracle against the REFERENCE's recorded answers (build container only: needs
/root/reference).  For a deterministic sample of the reference's COD10K images, apply the reference's preprocessing
(models/region_graph/extract_rg_embeddings.py:260-276: PIL convert('RGB').resize((256,256)), ToTensor / Normalize /
de-normalise / clip, then create_region_graph's (image*255).astype(uint8), :143), run oracle.slic.slic and compare the
number of regions with embedding_summary.json: images[name].num_nodes (what the reference's own skimage run produced).

Writes tests/golden/slic_counts.json = {"rows": [[name, expected, got], ...], "exact": n, ...} and
tests/golden/slic_images_u8.npz = the uint8 inputs of the first --fixture images (so the GPU box, which has no
/root/reference, can run cm_slic against the same recorded answers).

    python tools/check_slic_counts.py --n 240 --fixture 64
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def reference_uint8(path):
    """extract_embeddings_from_image's preprocessing (:260-276) + create_region_graph's uint8 cast (:143)."""
    from PIL import Image
    from torchvision import transforms
    image = Image.open(path).convert("RGB").resize((256, 256))
    t = transforms.Compose([transforms.ToTensor(), transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])(image)
    a = t.numpy().transpose(1, 2, 0)
    img = np.clip(a * np.array([0.229, 0.224, 0.225]) + np.array([0.485, 0.456, 0.406]), 0, 1)
    return img, (img * 255).astype(np.uint8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=240)
    ap.add_argument("--fixture", type=int, default=64)
    args = ap.parse_args()
    from oracle import slic as oslic
    summ = json.load(open(f"{REF}/models/region_graph/rg_embeddings/embedding_summary.json"))["images"]
    names = sorted(summ)
    stride = len(names) // args.n
    sample = names[::stride][:args.n]           # evenly over the sorted list: every super-class of COD10K
    rows, imgs = [], []
    for i, n in enumerate(sample):
        _, u8 = reference_uint8(f"{REF}/data/COD10K/images/{n}")
        got = int(len(np.unique(oslic.slic(u8))))
        rows.append([n, int(summ[n]["num_nodes"]), got])
        if i < args.fixture:
            imgs.append(u8)
    err = np.array([r[2] - r[1] for r in rows])
    hist = {str(int(k)): int((err == k).sum()) for k in np.unique(err)}
    out = {"source": "embedding_summary.json images[*].num_nodes vs oracle.slic.slic", "n": len(rows), "exact": int((err == 0).sum()),
           "within_1": int((np.abs(err) <= 1).sum()), "error_histogram": hist, "fixture_images": len(imgs), "rows": rows}
    with open(os.path.join(ROOT, "tests", "golden", "slic_counts.json"), "w") as f:
        json.dump(out, f, indent=0)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "slic_images_u8.npz"), images=np.stack(imgs),
                        names=np.array([r[0] for r in rows[:len(imgs)]]), expected=np.array([r[1] for r in rows[:len(imgs)]]))
    print(json.dumps({k: v for k, v in out.items() if k != "rows"}))


if __name__ == "__main__":
    main()
