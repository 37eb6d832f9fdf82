// Stand-in for <opencv2/imgproc.hpp> (TEST INFRASTRUCTURE, see core.hpp in this directory).
// cv::undistort (camera.hpp:54-58) is upstream of the hot path; the oracle harness feeds frames that are already
// undistorted, so this identity stub is never on a compared path.
#pragma once
#include <opencv2/core.hpp>
namespace cv {
inline void undistort(const Mat& in, Mat& out, const Mat&, const Mat&) { out = in.clone(); }
}
