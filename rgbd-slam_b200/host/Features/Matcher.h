// Matcher — same public interface as the reference's Features/Matcher.h:13-39.
// match() runs on the GPU (rgbdfe_match_knn2: K7 Hamming kNN k=2 + K8 ratio / unique-train /
// outlier / valid-depth filters), projectionMatch() too (rgbdfe_projection_match, scope row N2).
// drawMatches (HighGUI) is outside the path and not provided.
#ifndef MATCHER_H
#define MATCHER_H

#include "../System/Macros.h"
#include "../cvcompat.h"
#include <memory>
#include <tuple>
#include <vector>

class Frame;

class Matcher {
public:
    SMART_POINTER_TYPEDEFS(Matcher);

public:
    Matcher(float nnratio = 0.6);

    int match(std::shared_ptr<Frame> ref, std::shared_ptr<Frame> cur, std::vector<cv::DMatch>& vMatches12,
        const bool discardOutliers = true);

    int projectionMatch(std::shared_ptr<Frame> F1, std::shared_ptr<Frame> F2, std::vector<cv::DMatch>& vMatches12, const float th = 5.0f);

    // Computes the Hamming distance between two descriptors
    static double descriptorDistance(const cv::Mat& a, const cv::Mat& b);

    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;

protected:
    float mfNNratio;
};

#endif // ORBMATCHER_H
