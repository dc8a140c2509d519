// Mock of the PODs of reference src/features/FeatureExtractor.h:15-38 (same names and members).
#pragma once
#include <Eigen/Dense>
#include <opencv2/opencv.hpp>
#include <vector>
namespace slam {
class MapPoint;
struct ExtractedFeatures {
    std::vector<cv::KeyPoint> keypoints;
    cv::Mat descriptors;
};
struct FeatureMatch {
    FeatureMatch(int train_index, int query_index) : train_index(train_index), query_index(query_index) {}
    size_t train_index;
    size_t query_index;
};
struct MapPointMatch {
    MapPoint& point;
    size_t keypoint_index;
};
} // namespace slam
