# oracle/ref_paper.sed -- TEST INFRASTRUCTURE. Paper-mode hooks (SURVEY.md 8(f) N4) applied to the reference's rpf.cpp
# on top of ref_defined.sed (sed -n -f ref_paper.sed -f ref_defined.sed; addresses are reference line numbers).
# The 3-sigma threshold of rpf.cpp:579 goes through ref_paper_thresholds (ref_shim.cpp): 3*sigma unless the
# low-variance guard / world-position factor are set. Epsilon enters through REF_RATIO, which ref_defined.sed
# already puts at rpf.cpp:464,465,470.
579s|multiplyArray(pixelFstdDevMatrix\[pixel.x\]\[pixel.y\], 3)|::pbrt::ref_paper_thresholds(pixelFstdDevMatrix[pixel.x][pixel.y])|
