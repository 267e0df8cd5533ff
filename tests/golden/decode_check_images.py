"""Decodes the reference's own golden render check/lowres/03_volume_720_256.jpg (the one test
scene whose assets all ship) into tests/golden/check_03_volume_720_256.npy (uint8 sRGB,
[300, 720, 3]) so that the GPU box, which has no /root/reference, can compare against it.
Run where /root/reference exists:  python tests/golden/decode_check_images.py"""
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/check/lowres/03_volume_720_256.jpg"

if __name__ == "__main__":
    img = np.asarray(Image.open(SRC).convert("RGB"))
    print(img.shape, img.dtype, img.mean())
    np.save(os.path.join(HERE, "check_03_volume_720_256.npy"), img)
