// Solver base (reference: Solver/Solver.cpp:6-11).
#include "Solver.h"
#ifdef RGBDFE_WITH_OPENCV
#include "Core/Frame.h"
#else
#include "../Core/Frame.h"
#endif

using namespace std;

Solver::Solver(const Frame::Ptr F1, Frame::Ptr F2, const vector<cv::DMatch>& matches)
    : mF1(F1)
    , mF2(F2)
    , mMatches(matches)
{
}
