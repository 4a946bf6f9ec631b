// consumers.h — drop-in declarations of the reference entry points that consume the radius search and the
// range image next to the PCA stage, with the reference's names and signatures:
//   lidarOdometry::lidarOdometryClass::getRangeAndptIdImage / calculateSmoothness  (include/lidarOdometry.h:31-33)
//   DynaTools::clusterDBSCAN                                                       (include/dbscanCLuster.h:256-260)
//   buildingFacadeExtractor::meanShiftClustering / makeSuperVoxels                 (include/buildingFacadeExtraction.h)
//   preprocess::txt2pc / xyz2pc                                                    (src/smoothing.cpp:39-125)
// Bodies (consumers.cpp) forward to the CUDA path behind include/pcp.h; there is no CPU path.  Differences a
// caller can observe: the bool methods really return true/false (the reference falls off their ends), nothing
// is printed to std::cout, and where the reference indexes outside its 16 x 900 matrices the point is skipped.
#ifndef PCP_HOST_CONSUMERS_H_
#define PCP_HOST_CONSUMERS_H_

#include <string>
#include <vector>

#include "pcaAnalysis.h"  // pcXYZIptr / pcXYZI and the PCL / Eigen types (real or compat)

namespace lidarOdometry {
class lidarOdometryClass {
 public:
  bool getRangeAndptIdImage(pcXYZIptr inputCloud, Eigen::MatrixXf& rangeImg, Eigen::MatrixXf& ptIdImg);
  bool calculateSmoothness(Eigen::MatrixXf rangeImg, Eigen::MatrixXf& ptTypeImg);
};
}  // namespace lidarOdometry

namespace DynaTools {
int clusterDBSCAN(pcXYZIptr inCloud, double r, int minptNum, pcXYZIptr pcClusterCentos, pcXYZIptr pcClusterMinpts,
                  pcXYZIptr pcClusterMaxpts);
}

// preprocess::txt2pc (include/smoothing.h, src/smoothing.cpp:39-81): "x y z intensity" text file -> cloud (appended)
class preprocess {
 public:
  bool txt2pc(std::string txtpath, pcXYZI& origiCloud);
  bool xyz2pc(std::string txtpath, pcXYZI& origiCloud);  // src/smoothing.cpp:83-125
};

class buildingFacadeExtractor {
 public:
  struct voxel {  // include/buildingFacadeExtraction.h:24-30
    int id;
    int ptNums;
    std::vector<int> ptIndices;
    pcl::PointXYZI centroid;
  };
  bool meanShiftClustering(pcXYZIptr inputcloud, int iterTimes);
  // src/buildingFacadeExtraction.cpp:207-223: the PCA of every voxel about its centre, voxels with planeStruc > 0.6 are
  // appended to superVoxels.  ONE batched pcp_segment_pca call for all voxels (the reference: one calculatePCAofPoint
  // per voxel).  `features` (optional, not in the reference's signature) receives the record of every voxel.
  bool makeSuperVoxels(voxel* voxelArrptr, int voxelNums, pcXYZIptr inputCloud, std::vector<voxel>& superVoxels,
                       std::vector<pcaAnalysist::pcaFeature>* features = nullptr);
};

#endif
