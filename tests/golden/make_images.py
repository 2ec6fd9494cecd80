"""Writes tests/golden/images: one small picture in every EXR compression / pixel type the host reader supports plus
a Radiance .hdr, each written AND read back by OpenCV's codecs (the independent checker); the read-back floats are
stored next to each file as <file>.npy.  Run once here (needs cv2 with OpenEXR); the files are committed."""
import os

os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
import cv2
import numpy as np

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "images")
os.makedirs(OUT, exist_ok=True)
H, W = 35, 44
rng = np.random.default_rng(2024)
yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
img = np.stack([np.sin(xx * 0.2) * 0.5 + 0.5, (yy / (H - 1)) ** 2 * 30.0, np.cos(xx * 0.1 + yy * 0.13) * 2.0], -1).astype(np.float32)
img += rng.random((H, W, 3), dtype=np.float32) * 0.125
img[5, 7] = (4096.0, 0.0, 6e-5)
np.save(os.path.join(OUT, "expected.npy"), img)
for comp in ["NO", "RLE", "ZIPS", "ZIP", "PIZ", "PXR24"]:
    for ptype in ["HALF", "FLOAT"]:
        f = os.path.join(OUT, f"{comp.lower()}_{ptype.lower()}.exr")
        assert cv2.imwrite(f, img[:, :, ::-1], [cv2.IMWRITE_EXR_TYPE, getattr(cv2, "IMWRITE_EXR_TYPE_" + ptype),
                                                cv2.IMWRITE_EXR_COMPRESSION, getattr(cv2, "IMWRITE_EXR_COMPRESSION_" + comp)])
        back = np.ascontiguousarray(cv2.imread(f, cv2.IMREAD_UNCHANGED)[:, :, ::-1], np.float32)
        np.save(f + ".npy", back)
f = os.path.join(OUT, "radiance.hdr")
assert cv2.imwrite(f, img[:, :, ::-1])
np.save(f + ".npy", np.ascontiguousarray(cv2.imread(f, cv2.IMREAD_UNCHANGED)[:, :, ::-1], np.float32))
print(sorted(os.listdir(OUT)))
