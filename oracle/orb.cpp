// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// CPU restatement of the reference's ORB extractor:
//   Features/ORBextractor.cpp:16-41    IC_Angle
//   Features/ORBextractor.cpp:45-87    computeOrbDescriptor
//   Features/ORBextractor.cpp:348-406  constructor tables
//   Features/ORBextractor.cpp:414-611  DistributeOctTree
//   Features/ORBextractor.cpp:613-695  ComputeKeyPointsOctTree
//   Features/ORBextractor.cpp:706-766  operator()
//   Features/ORBextractor.cpp:773-797  ComputePyramid
//   Features/ExtractorNode.cpp:5-57    divideNode
// Determinism pin (the reference is address-dependent here): the
// quadtree's `sort(pair<int, ExtractorNode*>)` breaks size ties on the node
// pointer; the oracle breaks them on node creation order (later = greater).
#include "brief_pattern.h"
#include "oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <list>

namespace orc {

static const int PATCH_SIZE = 31;
static const int HALF_PATCH_SIZE = 15;
static const int EDGE_THRESHOLD = 19;

OrbExtractor::OrbExtractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures)
    , scaleFactor(_scaleFactor)
    , nlevels(_nlevels)
    , iniThFAST(_iniThFAST)
    , minThFAST(_minThFAST)
{
    mvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f;
    mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor); // float * double
        mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
        mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
    levels.resize(nlevels);

    mnFeaturesPerLevel.resize(nlevels);
    float factor = (float)(1.0f / scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        mnFeaturesPerLevel[level] = cv_round(nDesired);
        sum += mnFeaturesPerLevel[level];
        nDesired *= factor;
    }
    mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);

    umax.resize(HALF_PATCH_SIZE + 1);
    int v, v0, vmax = cv_floor(HALF_PATCH_SIZE * sqrtf(2.f) / 2 + 1);
    int vmin = (int)ceilf(HALF_PATCH_SIZE * sqrtf(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v)
        umax[v] = cv_round_d(sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1])
            ++v0;
        umax[v] = v0;
        ++v0;
    }
}

void OrbExtractor::compute_pyramid(const uint8_t* img, int w, int h, int stride)
{
    for (int level = 0; level < nlevels; ++level) {
        float scale = mvInvScaleFactor[level];
        OrbLevel& L = levels[level];
        L.w = cv_round((float)w * scale);
        L.h = cv_round((float)h * scale);
        L.stride = L.w + EDGE_THRESHOLD * 2;
        L.padded.assign((size_t)L.stride * (L.h + EDGE_THRESHOLD * 2), 0);
        L.blurred.clear();
        std::vector<uint8_t> tmp((size_t)L.w * L.h);
        if (level != 0) {
            const OrbLevel& P = levels[level - 1];
            resize_linear_u8(P.image(), P.w, P.h, P.stride, tmp.data(), L.w, L.h, L.w);
            border_reflect101(tmp.data(), L.w, L.h, L.w, L.padded.data(), L.stride, EDGE_THRESHOLD);
        } else {
            border_reflect101(img, w, h, stride, L.padded.data(), L.stride, EDGE_THRESHOLD);
        }
    }
}

namespace {
    struct Node {
        std::vector<KeyPoint> keys;
        int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
        std::list<Node>::iterator lit;
        bool noMore = false;
        int seq = 0;

        void divide(Node& n1, Node& n2, Node& n3, Node& n4) const
        {
            const int halfX = (int)ceilf(static_cast<float>(URx - ULx) / 2);
            const int halfY = (int)ceilf(static_cast<float>(BRy - ULy) / 2);
            n1.ULx = ULx, n1.ULy = ULy;
            n1.URx = ULx + halfX, n1.URy = ULy;
            n1.BLx = ULx, n1.BLy = ULy + halfY;
            n1.BRx = ULx + halfX, n1.BRy = ULy + halfY;
            n2.ULx = n1.URx, n2.ULy = n1.URy;
            n2.URx = URx, n2.URy = URy;
            n2.BLx = n1.BRx, n2.BLy = n1.BRy;
            n2.BRx = URx, n2.BRy = ULy + halfY;
            n3.ULx = n1.BLx, n3.ULy = n1.BLy;
            n3.URx = n1.BRx, n3.URy = n1.BRy;
            n3.BLx = BLx, n3.BLy = BLy;
            n3.BRx = n1.BRx, n3.BRy = BLy;
            n4.ULx = n3.URx, n4.ULy = n3.URy;
            n4.URx = n2.BRx, n4.URy = n2.BRy;
            n4.BLx = n3.BRx, n4.BLy = n3.BRy;
            n4.BRx = BRx, n4.BRy = BRy;
            for (const KeyPoint& kp : keys) {
                if (kp.x < n1.URx) {
                    if (kp.y < n1.BRy)
                        n1.keys.push_back(kp);
                    else
                        n3.keys.push_back(kp);
                } else if (kp.y < n1.BRy)
                    n2.keys.push_back(kp);
                else
                    n4.keys.push_back(kp);
            }
            n1.noMore = n1.keys.size() == 1;
            n2.noMore = n2.keys.size() == 1;
            n3.noMore = n3.keys.size() == 1;
            n4.noMore = n4.keys.size() == 1;
        }
    };
}

std::vector<KeyPoint> OrbExtractor::distribute_quadtree(const std::vector<KeyPoint>& keys, int minX, int maxX, int minY, int maxY, int N)
{
    const int nIni = (int)roundf(static_cast<float>(maxX - minX) / (maxY - minY));
    const float hX = static_cast<float>(maxX - minX) / nIni;
    int seq = 0;

    std::list<Node> nodes;
    std::vector<Node*> ini(nIni);
    for (int i = 0; i < nIni; i++) {
        Node ni;
        ni.ULx = (int)(hX * static_cast<float>(i)), ni.ULy = 0;
        ni.URx = (int)(hX * static_cast<float>(i + 1)), ni.URy = 0;
        ni.BLx = ni.ULx, ni.BLy = maxY - minY;
        ni.BRx = ni.URx, ni.BRy = maxY - minY;
        ni.seq = seq++;
        nodes.push_back(ni);
        ini[i] = &nodes.back();
    }
    for (const KeyPoint& kp : keys)
        ini[(size_t)(kp.x / hX)]->keys.push_back(kp);

    auto lit = nodes.begin();
    while (lit != nodes.end()) {
        if (lit->keys.size() == 1) {
            lit->noMore = true;
            lit++;
        } else if (lit->keys.empty())
            lit = nodes.erase(lit);
        else
            lit++;
    }

    bool finish = false;
    typedef std::pair<int, int> SizeSeq; // (size, creation seq)
    std::vector<std::pair<SizeSeq, Node*>> sizeAndNode;

    auto push_children = [&](Node* c[4], int* nToExpand) {
        for (int k = 0; k < 4; k++) {
            if (c[k]->keys.size() > 0) {
                c[k]->seq = seq++;
                nodes.push_front(*c[k]);
                if (c[k]->keys.size() > 1) {
                    if (nToExpand)
                        (*nToExpand)++;
                    sizeAndNode.push_back(std::make_pair(SizeSeq((int)c[k]->keys.size(), nodes.front().seq), &nodes.front()));
                    nodes.front().lit = nodes.begin();
                }
            }
        }
    };

    while (!finish) {
        int prevSize = (int)nodes.size();
        lit = nodes.begin();
        int nToExpand = 0;
        sizeAndNode.clear();

        while (lit != nodes.end()) {
            if (lit->noMore) {
                lit++;
                continue;
            }
            Node n1, n2, n3, n4;
            lit->divide(n1, n2, n3, n4);
            Node* c[4] = { &n1, &n2, &n3, &n4 };
            push_children(c, &nToExpand);
            lit = nodes.erase(lit);
        }

        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finish = true;
        } else if (((int)nodes.size() + nToExpand * 3) > N) {
            while (!finish) {
                prevSize = (int)nodes.size();
                std::vector<std::pair<SizeSeq, Node*>> prev = sizeAndNode;
                sizeAndNode.clear();
                std::sort(prev.begin(), prev.end(),
                    [](const std::pair<SizeSeq, Node*>& a, const std::pair<SizeSeq, Node*>& b) { return a.first < b.first; });
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    Node n1, n2, n3, n4;
                    prev[j].second->divide(n1, n2, n3, n4);
                    Node* c[4] = { &n1, &n2, &n3, &n4 };
                    push_children(c, nullptr);
                    nodes.erase(prev[j].second->lit);
                    if ((int)nodes.size() >= N)
                        break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize)
                    finish = true;
            }
        }
    }

    std::vector<KeyPoint> result;
    result.reserve(nodes.size());
    for (auto it = nodes.begin(); it != nodes.end(); it++) {
        const std::vector<KeyPoint>& nk = it->keys;
        const KeyPoint* best = &nk[0];
        float maxResponse = best->response;
        for (size_t k = 1; k < nk.size(); k++) {
            if (nk[k].response > maxResponse) {
                best = &nk[k];
                maxResponse = nk[k].response;
            }
        }
        result.push_back(*best);
    }
    return result;
}

float OrbExtractor::ic_angle(const OrbLevel& L, float px, float py) const
{
    int m_01 = 0, m_10 = 0;
    const int step = L.stride;
    const uint8_t* center = L.image() + cv_round(py) * step + cv_round(px);
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u)
        m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0;
        int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return fast_atan2((float)m_01, (float)m_10);
}

void OrbExtractor::compute_keypoints()
{
    const float W = 30;
    for (int level = 0; level < nlevels; ++level) {
        OrbLevel& L = levels[level];
        const int minBorderX = EDGE_THRESHOLD - 3;
        const int minBorderY = minBorderX;
        const int maxBorderX = L.w - EDGE_THRESHOLD + 3;
        const int maxBorderY = L.h - EDGE_THRESHOLD + 3;

        std::vector<KeyPoint>& cand = L.candidates;
        cand.clear();

        const float width = (float)(maxBorderX - minBorderX);
        const float height = (float)(maxBorderY - minBorderY);
        const int nCols = (int)(width / W);
        const int nRows = (int)(height / W);
        const int wCell = (int)ceilf(width / nCols);
        const int hCell = (int)ceilf(height / nRows);

        std::vector<int> buf(3 * 64 * 64);
        StageTimer* t_fast = new StageTimer(ST_FAST);
        for (int i = 0; i < nRows; i++) {
            const float iniY = (float)(minBorderY + i * hCell);
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBorderY - 3)
                continue;
            if (maxY > maxBorderY)
                maxY = (float)maxBorderY;
            for (int j = 0; j < nCols; j++) {
                const float iniX = (float)(minBorderX + j * wCell);
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBorderX - 6)
                    continue;
                if (maxX > maxBorderX)
                    maxX = (float)maxBorderX;

                const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
                const uint8_t* roi = L.image() + (size_t)y0 * L.stride + x0;
                if ((size_t)cw * ch * 3 > buf.size())
                    buf.resize((size_t)cw * ch * 3);
                int n = fast9(roi, cw, ch, L.stride, iniThFAST, 1, buf.data(), cw * ch);
                if (n == 0)
                    n = fast9(roi, cw, ch, L.stride, minThFAST, 1, buf.data(), cw * ch);
                for (int k = 0; k < n; k++) {
                    KeyPoint kp;
                    kp.x = (float)buf[k * 3] + j * wCell;
                    kp.y = (float)buf[k * 3 + 1] + i * hCell;
                    kp.size = 7.f;
                    kp.angle = -1.f;
                    kp.response = (float)buf[k * 3 + 2];
                    kp.octave = 0;
                    kp.class_id = -1;
                    cand.push_back(kp);
                }
            }
        }

        delete t_fast;
        StageTimer t_qt(ST_QUADTREE);
        L.keypoints = distribute_quadtree(cand, minBorderX, maxBorderX, minBorderY, maxBorderY, mnFeaturesPerLevel[level]);
        const int scaledPatchSize = (int)(PATCH_SIZE * mvScaleFactor[level]);
        for (KeyPoint& kp : L.keypoints) {
            kp.x += minBorderX;
            kp.y += minBorderY;
            kp.octave = level;
            kp.size = (float)scaledPatchSize;
        }
    }
    StageTimer t_ang(ST_ANGLE);
    for (int level = 0; level < nlevels; ++level)
        for (KeyPoint& kp : levels[level].keypoints)
            kp.angle = ic_angle(levels[level], kp.x, kp.y);
}

void OrbExtractor::describe(const uint8_t* img, int stride, const KeyPoint& kp, uint8_t* desc)
{
    const float factorPI = (float)(M_PI / 180.f);
    float angle = (float)kp.angle * factorPI;
    float a = cosf_glibc(angle), b = sinf_glibc(angle);
    const uint8_t* center = img + cv_round(kp.y) * stride + cv_round(kp.x);
    const signed char* pat = kBriefPattern;
    for (int i = 0; i < 32; ++i, pat += 32) {
        int val = 0;
        for (int k = 0; k < 8; k++) {
            const signed char* p0 = pat + k * 4; // (x0,y0,x1,y1) of test k in this byte
            int t0 = center[cv_round(p0[0] * b + p0[1] * a) * stride + cv_round(p0[0] * a - p0[1] * b)];
            int t1 = center[cv_round(p0[2] * b + p0[3] * a) * stride + cv_round(p0[2] * a - p0[3] * b)];
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

void OrbExtractor::extract(const uint8_t* img, int w, int h, int stride, std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc)
{
    kps.clear();
    desc.clear();
    if (!img || w <= 0 || h <= 0)
        return;
    {
        StageTimer t(ST_PYRAMID);
        compute_pyramid(img, w, h, stride);
    }
    compute_keypoints();

    int total = 0;
    for (int level = 0; level < nlevels; ++level)
        total += (int)levels[level].keypoints.size();
    desc.resize((size_t)total * 32);
    kps.reserve(total);
    int offset = 0;
    for (int level = 0; level < nlevels; ++level) {
        OrbLevel& L = levels[level];
        int n = (int)L.keypoints.size();
        if (n == 0)
            continue;
        // clone() of the un-padded ROI, then GaussianBlur with REFLECT_101
        L.blurred.resize((size_t)L.w * L.h);
        {
            StageTimer t(ST_BLUR);
            gauss7_u8(L.image(), L.w, L.h, L.stride, L.blurred.data(), L.w);
        }
        StageTimer t(ST_DESCRIBE);
        for (int i = 0; i < n; i++)
            describe(L.blurred.data(), L.w, L.keypoints[i], desc.data() + (size_t)(offset + i) * 32);
        offset += n;
        float scale = mvScaleFactor[level];
        for (int i = 0; i < n; i++) {
            KeyPoint kp = L.keypoints[i];
            if (level != 0) {
                kp.x *= scale;
                kp.y *= scale;
            }
            kps.push_back(kp);
        }
    }
}

} // namespace orc
