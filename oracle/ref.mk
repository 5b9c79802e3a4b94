# oracle/ref.mk — builds oracle/_ref/liborb_ref.so: the reference's OWN src/geometry/fextractor.cpp,
# compiled unmodified from where it lies under $(REF), against the OpenCV stand-in in oracle/ref_compat/
# (types + the cv2-pinned primitive restatements of orb_oracle.cpp).  Test infrastructure only.
# Outputs go to oracle/_ref/ only (git-ignored, travels to the GPU box with gpurun).
# Not the reference's build system: the CMake project needs OpenCV 4.2, Eigen, Sophus, Pangolin, g2o, DBoW3.
#   make -C oracle -f ref.mk            (REF=/root/reference by default)
REF ?= /root/reference
CXX ?= g++
# -O3 without -march=native: reference CMakeLists.txt:14-15.  -ffp-contract=off is a no-op on baseline x86-64
# (no FMA) and keeps the float index math of computeOrbDescriptor uncontracted elsewhere.
CXXFLAGS ?= -O3 -std=c++14 -fPIC -pthread -ffp-contract=off -w
OUT := _ref/liborb_ref.so
SRC := $(REF)/src/geometry/fextractor.cpp ref_compat/cv_compat.cpp ref_compat/ref_driver.cpp orb_oracle.cpp

all: $(OUT)

$(OUT): $(SRC) ref_compat/cv_compat.hpp ../vi_slam_b200/csrc/orb_pattern.inc
	mkdir -p _ref
	$(CXX) $(CXXFLAGS) -Iref_compat -I$(REF)/include -shared -Wl,-Bsymbolic -o $@ \
	    $(REF)/src/geometry/fextractor.cpp ref_compat/cv_compat.cpp ref_compat/ref_driver.cpp \
	    -x c++ -std=c++17 orb_oracle.cpp
	sha256sum $(REF)/src/geometry/fextractor.cpp $(REF)/include/vi_slam/geometry/fextractor.h > _ref/SOURCES.sha256

clean:
	rm -rf _ref
