"""Writes the small EXR fixtures under tests/golden/ with OpenCV's OpenEXR codec (an implementation independent of the two
readers under test: spt/exrio.py and host/exr.cpp).  Run once in the build container:
    OPENCV_IO_ENABLE_OPENEXR=1 python tools/make_golden_exr.py
Each file holds the same 24x12 lat-long sky (smooth gradient + one very bright texel); env_golden.npy is the f32 source."""
import os

os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
import cv2  # noqa: E402
import numpy as np  # noqa: E402

G = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
h, w = 12, 24
y, x = np.mgrid[0:h, 0:w]
img = np.stack([0.25 + 0.5 * y / h, 0.5 + 0.25 * np.cos(x / w * 2 * np.pi), 1.0 - 0.5 * y / h], -1).astype(np.float32)
img = np.round(img * 64) / 64  # exactly representable in half, so every file decodes to the same numbers
img[3, 17] = [1000.0, 944.0, 848.0]
np.save(os.path.join(G, "env_golden.npy"), img)
for cname, cflag in (("none", 0), ("rle", 1), ("zips", 2), ("zip", 3)):
    for tname, tflag in (("float", cv2.IMWRITE_EXR_TYPE_FLOAT), ("half", cv2.IMWRITE_EXR_TYPE_HALF)):
        fn = os.path.join(G, f"env_{cname}_{tname}.exr")
        assert cv2.imwrite(fn, img[..., ::-1], [cv2.IMWRITE_EXR_TYPE, tflag, cv2.IMWRITE_EXR_COMPRESSION, cflag])
        print(fn, os.path.getsize(fn))
assert cv2.imwrite(os.path.join(G, "env_piz_half.exr"), img[..., ::-1], [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_HALF, cv2.IMWRITE_EXR_COMPRESSION, 4])
