"""Generate tests/golden/pipeline_aug.npz from the UNMODIFIED reference input-pipeline functions (build container only):
MoCE-IR-main/src/utils/image_utils.py data_augmentation and data/degradation_utils.py Degradation._add_gaussian_noise,
on a small synthetic uint8 image, for every augmentation mode and the three noise levels of the reference."""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/MoCE-IR-main/src")
from utils import image_utils as IU  # noqa: E402  (the reference)
from data import degradation_utils as DU  # noqa: E402  (the reference)

GOLD = os.path.join(ROOT, "tests", "golden")
P = 16


def main():
    rng = np.random.RandomState(123)
    img = rng.randint(0, 256, size=(40, 52, 3)).astype(np.uint8)
    D = DU.Degradation(types.SimpleNamespace(patch_size=P))
    rec = {"img": img, "offs": np.array([[3, 7], [24, 36], [0, 0], [11, 20]], dtype=np.int32)}
    for i, (oy, ox) in enumerate(rec["offs"]):
        patch = img[oy:oy + P, ox:ox + P]
        for mode in range(8):
            aug = np.ascontiguousarray(IU.data_augmentation(patch, mode)) if mode else patch
            rec[f"aug_{i}_{mode}"] = aug
            for sigma in (15, 25, 50):
                np.random.seed(1000 + 10 * mode + i)
                noisy, _ = D._add_gaussian_noise(aug, sigma=sigma)
                rec[f"noisy_{i}_{mode}_{sigma}"] = noisy
    np.savez_compressed(os.path.join(GOLD, "pipeline_aug.npz"), **rec)
    print("pipeline_aug", len(rec))


if __name__ == "__main__":
    main()
