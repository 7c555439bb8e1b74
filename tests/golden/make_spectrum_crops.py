#!/usr/bin/env python
"""More of the reference's own DFT_image pairs as small fixtures (run in the build container, /root/reference mounted):

    python tests/golden/make_spectrum_crops.py

For six blur radii the input PNG (test_images/spectrum_analysis/blur/spatial domain/blur_radiusR.png) is copied as
spectrum_in_blurR.png and rows [480, 544) of the reference's spectrum image (.../time domain/radius_R.png: the 64
rows around the DC term, where the spectrum has structure) are stored as spectrum_out_blurR_rows480_544.png.
The GPU test compares blur_b200_dft_image on the full input with these rows (tests/test_gpu_parity.py).
"""
import os
import shutil

import cv2

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/test_images/spectrum_analysis/blur"
RADII = (10, 30, 50, 70, 90, 100)
R0, R1 = 480, 544


def main() -> None:
    for r in RADII:
        shutil.copyfile(os.path.join(SRC, "spatial domain", f"blur_radius{r}.png"), os.path.join(HERE, f"spectrum_in_blur{r}.png"))
        os.chmod(os.path.join(HERE, f"spectrum_in_blur{r}.png"), 0o644)
        out = cv2.imread(os.path.join(SRC, "time domain", f"radius_{r}.png"), cv2.IMREAD_COLOR)
        assert out is not None and out.shape[0] >= R1, out.shape
        cv2.imwrite(os.path.join(HERE, f"spectrum_out_blur{r}_rows{R0}_{R1}.png"), out[R0:R1], [cv2.IMWRITE_PNG_COMPRESSION, 9])
        print(r, out.shape)


if __name__ == "__main__":
    main()
