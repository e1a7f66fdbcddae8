# oracle/ref_paper.sed -- TEST INFRASTRUCTURE. Paper-mode hooks (SURVEY.md 8(f) N4) applied to the reference's rpf.cpp
# on top of ref_defined.sed (sed -n -f ref_paper.sed -f ref_defined.sed; addresses are reference line numbers).
# The 3-sigma threshold of rpf.cpp:579 goes through ref_paper_thresholds (ref_shim.cpp): 3*sigma unless the
# low-variance guard / world-position factor are set. Epsilon enters through REF_RATIO, which ref_defined.sed
# already puts at rpf.cpp:464,465,470.
579s|multiplyArray(pixelFstdDevMatrix\[pixel.x\]\[pixel.y\], 3)|::pbrt::ref_paper_thresholds(pixelFstdDevMatrix[pixel.x][pixel.y])|
# Random neighbour subsampling (rpf_set_subsampling): a sample counter for the range-for of rpf.cpp:573 and the keep
# decision of include/rpf_subsample.h joined to the 3-sigma test at rpf.cpp:581.
573s|for (const SampleData \&sf|int ref_si_ = -1; for (const SampleData \&sf|
581s|if (within3StdDevs) {|++ref_si_; if (within3StdDevs \&\& ::pbrt::ref_paper_keep(box_size, pixel.x, pixel.y, xn, yn, ref_si_)) {|
