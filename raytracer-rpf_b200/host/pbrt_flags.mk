# Compile / link flags of the reference's own Release build (CMakeLists.txt:72-99,105-288), as `ninja -t commands pbrt_exe` prints
# them in $(REFBUILD). Shared by host/Makefile (the product's plugin binary) and oracle/Makefile (the reference dump hook).
PBRT_DEFS = -D PBRT_HAVE_ALLOCA_H -D PBRT_HAVE_MEMORY_H -D PBRT_HAVE_HEX_FP_CONSTANTS -D PBRT_HAVE_BINARY_CONSTANTS \
  -D PBRT_HAVE_CONSTEXPR -D PBRT_CONSTEXPR=constexpr -D PBRT_HAVE_ALIGNAS -D PBRT_HAVE_ALIGNOF -D PBRT_HAVE_ITIMER \
  -D PBRT_HAVE_NONPOD_IN_UNIONS -D PBRT_HAVE_MMAP '-D PBRT_NOINLINE=__attribute__((noinline))' -D PBRT_HAVE_POSIX_MEMALIGN \
  -D PBRT_THREAD_LOCAL=thread_local
PBRT_INC = -I$(REF)/src/ext/openexr/IlmBase/Imath -I$(REF)/src/ext/openexr/IlmBase/Half -I$(REF)/src/ext/openexr/IlmBase/Iex \
  -I$(REF)/src/ext/openexr/OpenEXR/IlmImf -I$(REFBUILD)/src/ext/openexr/IlmBase/config -I$(REFBUILD)/src/ext/openexr/OpenEXR/config \
  -I$(REF)/src/ext/glog/src -I$(REFBUILD)/src/ext/glog -I$(REF)/src/ext/ptex/src/ptex -I$(REF)/src -I$(REF)/src/core
# the reference's own flags for its sources (CMakeLists.txt:72-99, Release)
PBRT_CXXFLAGS = -std=gnu++11 -Wno-conversion-null -w -O3 -DNDEBUG $(PBRT_DEFS) $(PBRT_INC)
PBRT_LIBS = $(REFBUILD)/libpbrt.a $(REFBUILD)/src/ext/openexr/OpenEXR/IlmImf/libIlmImf.a $(REFBUILD)/src/ext/openexr/IlmBase/Imath/libImath.a \
  $(REFBUILD)/src/ext/openexr/IlmBase/Half/libHalf.a $(REFBUILD)/src/ext/glog/libglog.a $(REFBUILD)/src/ext/ptex/src/ptex/libPtex.a \
  $(REFBUILD)/src/ext/openexr/IlmBase/IlmThread/libIlmThread.a $(REFBUILD)/src/ext/openexr/IlmBase/Iex/libIex.a -lz -lpthread

