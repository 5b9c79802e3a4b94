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
FRAME_CPP := $(REF)/src/datastructures/frame.cpp
FMATCHER_CPP := $(REF)/src/geometry/fmatcher.cpp
PINHOLE_CPP := $(REF)/src/geometry/cameramodels/pinhole.cpp
DBOW3 := $(REF)/thirdparty/DBoW3/DBoW3/src
KB8_CPP := $(REF)/src/geometry/cameramodels/kannalabrandt8.cpp
SRC := $(REF)/src/geometry/fextractor.cpp ref_compat/cv_compat.cpp ref_compat/ref_driver.cpp ref_compat/ref_slices.cpp orb_oracle.cpp \
       $(FRAME_CPP) $(FMATCHER_CPP) $(PINHOLE_CPP) ref_compat/ref_dbow3.cpp $(DBOW3)/Vocabulary.cpp $(DBOW3)/DescManip.cpp \
       $(DBOW3)/BowVector.cpp $(DBOW3)/FeatureVector.cpp ref_compat/ref_kb8.cpp ref_compat/ref_matchfeatures.cpp $(KB8_CPP)

all: $(OUT)

# Function definitions of frame.cpp / fmatcher.cpp cut out by line range (those translation units need Eigen, Sophus, g2o,
# DBoW3 ... and cannot be compiled whole; ref_compat/ref_slices.cpp says what the slices are compiled against).
# slice <file> <first> <last> <name> <text the first line must contain>
define slice
	@sed -n '$(2)p' $(1) | grep -q -F '$(5)' || { echo "ref.mk: $(1):$(2) is not '$(5)' - the reference changed" >&2; exit 1; }
	@sed -n '$(3)p' $(1) | grep -q '^ *}' || { echo "ref.mk: $(1):$(3) does not close a function" >&2; exit 1; }
	sed -n '$(2),$(3)p' $(1) > _ref/gen/$(4)_$(2)_$(3).inc
endef

$(OUT): $(SRC) ref_compat/cv_compat.hpp ../vi_slam_b200/csrc/orb_pattern.inc ref.mk
	mkdir -p _ref/gen
	$(call slice,$(FRAME_CPP),386,414,frame,void Frame::AssignFeaturesToGrid)
	$(call slice,$(FRAME_CPP),678,744,frame,Frame::GetFeaturesInArea)
	$(call slice,$(FRAME_CPP),746,756,frame,bool Frame::PosInGrid)
	$(call slice,$(FRAME_CPP),823,997,frame,void Frame::ComputeStereoMatches)
	$(call slice,$(FRAME_CPP),758,791,frame,void Frame::UndistortKeyPoints)
	$(call slice,$(FRAME_CPP),793,821,frame,void Frame::ComputeImageBounds)
	$(call slice,$(FMATCHER_CPP),313,319,fmatcher,const int FMatcher::TH_HIGH = 100;)
	$(call slice,$(FMATCHER_CPP),983,1098,fmatcher,int FMatcher::SearchForInitialization)
	$(call slice,$(FMATCHER_CPP),2813,2854,fmatcher,void FMatcher::ComputeThreeMaxima)
	$(call slice,$(FMATCHER_CPP),2859,2875,fmatcher,int FMatcher::DescriptorDistance)
	$(call slice,$(FMATCHER_CPP),132,245,fmatcher,void matchFeatures)
	$(call slice,$(FMATCHER_CPP),247,266,fmatcher,void removeDuplicatedMatches)
	$(call slice,$(FMATCHER_CPP),2471,2687,fmatcher,const Frame &LastFrame)
	$(call slice,$(PINHOLE_CPP),13,16,pinhole,const cv::Point3f &p3D)
	$(call slice,$(PINHOLE_CPP),22,26,pinhole,const cv::Mat &m3D)
	$(call slice,$(DBOW3)/Vocabulary.cpp,754,820,Vocabulary,void Vocabulary::transform)
	$(call slice,$(DBOW3)/Vocabulary.cpp,838,876,Vocabulary,void Vocabulary::transform)
	$(call slice,$(DBOW3)/DescManip.cpp,92,130,DescManip,double DescManip::distance)
	$(call slice,$(KB8_CPP),11,27,kannalabrandt8,KannalaBrandt8::project)
	$(call slice,$(KB8_CPP),33,37,kannalabrandt8,KannalaBrandt8::project)
	$(call slice,$(KB8_CPP),70,74,kannalabrandt8,KannalaBrandt8::unprojectMat)
	$(call slice,$(KB8_CPP),82,109,kannalabrandt8,KannalaBrandt8::unproject)
	$(call slice,$(KB8_CPP),324,393,kannalabrandt8,KannalaBrandt8::TriangulateMatches)
	$(call slice,$(KB8_CPP),485,498,kannalabrandt8,void KannalaBrandt8::Triangulate)
	$(call slice,$(FRAME_CPP),1149,1189,frame,void Frame::ComputeStereoFishEyeMatches)
	@# -include ...: the vendored DBoW3 headers rely on transitive std includes that GCC 13's leaner headers no longer provide
	$(CXX) $(CXXFLAGS) -Iref_compat -I_ref -I$(REF)/include -shared -Wl,-Bsymbolic -o $@ \
	    $(REF)/src/geometry/fextractor.cpp ref_compat/cv_compat.cpp ref_compat/ref_driver.cpp ref_compat/ref_slices.cpp \
	    ref_compat/ref_kb8.cpp ref_compat/ref_matchfeatures.cpp -I$(DBOW3) -include cstdint -include iostream -include string ref_compat/ref_dbow3.cpp $(DBOW3)/BowVector.cpp $(DBOW3)/FeatureVector.cpp \
	    -x c++ -std=c++17 orb_oracle.cpp
	sha256sum $(REF)/src/geometry/fextractor.cpp $(REF)/include/vi_slam/geometry/fextractor.h $(FRAME_CPP) $(FMATCHER_CPP) $(PINHOLE_CPP) \
	    $(DBOW3)/Vocabulary.cpp $(DBOW3)/DescManip.cpp $(DBOW3)/BowVector.cpp $(DBOW3)/FeatureVector.cpp $(KB8_CPP) > _ref/SOURCES.sha256
	rm -rf _ref/gen          # the slices are build intermediates: only the library stays

clean:
	rm -rf _ref
