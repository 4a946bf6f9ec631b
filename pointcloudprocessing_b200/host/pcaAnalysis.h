// pcaAnalysis.h — drop-in declaration of the reference's `pcaAnalysist` class
// (reference: include/pcaAnalysis.h:14-60; typedefs include/smoothing.h:31-32).
//
// The public surface — the nested pcaFeature record and the four methods with their exact
// signatures — is what callers such as buildingFacadeExtractor::makeSuperVoxels
// (src/buildingFacadeExtraction.cpp:212-219) compile against, so it is kept as is; the bodies in
// pcaAnalysis.cpp forward to the CUDA path behind the C-ABI of include/pcp.h.  Differences a caller
// can observe: the methods now really return true/false (the reference falls off the end of
// non-void functions), and records of the leading below-gate points are zero-filled
// instead of uninitialised (Q4).
#ifndef PCP_HOST_PCAANALYSIS_H_
#define PCP_HOST_PCAANALYSIS_H_

#ifdef PCP_HAVE_PCL
#include "smoothing.h"   // the reference's own header: PCL types + pcXYZIptr
#include <Eigen/Dense>
#else
#include "compat/pcl_min.h"
typedef pcl::PointCloud<pcl::PointXYZI>::Ptr pcXYZIptr;
typedef pcl::PointCloud<pcl::PointXYZI> pcXYZI;
#endif

#include <math.h>
#include <vector>

class pcaAnalysist {
  struct lamda {  // eigenvalues, ascending: la1 <= la2 <= la3
    float la1;
    float la2;
    float la3;
  };
  struct vec {  // eigenvectors of la1, la2, la3
    Eigen::Vector3f v1;
    Eigen::Vector3f v2;
    Eigen::Vector3f v3;
  };

 public:
  struct pcaFeature {
    lamda lamds;
    vec vecs;
    int ptID;
    int ptNum;  // neighbours found, the point itself included
    float lineStruc;
    float planeStruc;
    float scatterStruc;
  };

  // Whole-cloud entry point (reference src/pcaAnalysis.cpp:11-42): radius search around every
  // point, PCA about the point when it has >= 13 neighbours, one record appended per point.
  bool calculatePCAofPointCloud(pcXYZIptr inputCloud, float radius, std::vector<pcaFeature>& pcaFeaofCloud);

  // One point with a caller-supplied neighbour list (src/pcaAnalysis.cpp:44-87).
  bool calculatePCAofPoint(pcXYZIptr inputCloud, int i, std::vector<int> neighborIndices, pcaFeature& ptFeature);

  // One segment about an arbitrary centre, no minimum-count gate (src/pcaAnalysis.cpp:89-121).
  bool calculatePCAofPoint(pcXYZIptr inputCloud, pcl::PointXYZI centro, std::vector<int> neighborIndices,
                           pcaAnalysist::pcaFeature& ptFeature);

  // Whole-cloud PCA frame and extents (src/pcaAnalysis.cpp:123-148; the reference discards the
  // result and dereferences a null cloud, SURVEY Q9 — here it computes and returns true).
  bool getMiniBoundingBox(pcXYZIptr inputCloud);
};

static_assert(sizeof(pcaAnalysist::pcaFeature) == 68, "pcaFeature must stay layout-identical to pcp_feature68");

#endif  // PCP_HOST_PCAANALYSIS_H_
