// ORACLE -- TEST INFRASTRUCTURE ONLY (see orbref_cv.h).  CPU restatement of the reference's
// Hamming matchers (ORB_SLAM2/src/ORBmatcher.cc) on flat arrays.
#include "orbref.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

// ref ORBmatcher::DescriptorDistance, ORBmatcher.cc:1650-1666
static inline int descriptor_distance(const uint8_t *a, const uint8_t *b) {
    const int32_t *pa = (const int32_t *)a, *pb = (const int32_t *)b;
    int dist = 0;
    for (int i = 0; i < 8; i++, pa++, pb++) {
        unsigned int v = *pa ^ *pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}
// same distance with the hardware popcount (used only by the timed CPU baseline)
static inline int descriptor_distance_popcnt(const uint8_t *a, const uint8_t *b) {
    const uint64_t *pa = (const uint64_t *)a, *pb = (const uint64_t *)b;
    return __builtin_popcountll(pa[0] ^ pb[0]) + __builtin_popcountll(pa[1] ^ pb[1]) + __builtin_popcountll(pa[2] ^ pb[2]) +
           __builtin_popcountll(pa[3] ^ pb[3]);
}

extern "C" {

// Brute-force best / second-best scan with the reference's update rule (ORBmatcher.cc:216-225):
// strict '<' on both, so the first row seen wins ties.  fast=1 uses the hardware popcount and
// `threads` workers over query ranges (CPU baseline); fast=0 is the literal SWAR loop.
int orbref_hamming_top2(const uint8_t *queries, int nq, const uint8_t *db, int64_t ndb, int64_t idx_base,
                        int32_t *best_dist, int64_t *best_idx, int32_t *second_dist, int64_t *second_idx, int fast,
                        int threads) {
    auto work = [&](int q0, int q1) {
        for (int q = q0; q < q1; q++) {
            const uint8_t *dq = queries + (size_t)q * 32;
            int d1 = 256, d2 = 256;
            int64_t i1 = -1, i2 = -1;
            for (int64_t r = 0; r < ndb; r++) {
                const int d = fast ? descriptor_distance_popcnt(dq, db + (size_t)r * 32) : descriptor_distance(dq, db + (size_t)r * 32);
                if (d < d1) { d2 = d1; i2 = i1; d1 = d; i1 = r; }
                else if (d < d2) { d2 = d; i2 = r; }
            }
            best_dist[q] = d1; second_dist[q] = d2;
            best_idx[q] = i1 < 0 ? -1 : i1 + idx_base;
            second_idx[q] = i2 < 0 ? -1 : i2 + idx_base;
        }
    };
    if (threads <= 1) { work(0, nq); return ORB_OK; }
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++) th.emplace_back(work, (int)((int64_t)nq * t / threads), (int)((int64_t)nq * (t + 1) / threads));
    for (auto &t : th) t.join();
    return ORB_OK;
}

// Relocalisation candidate score of a database made of keyframes of `kf_rows` consecutive rows: counts[k] = number of queries
// whose best row inside keyframe k passes the reference's acceptance for a (query, keyframe) pair, ORBmatcher.cc:216-231:
// bestDist1 <= TH_LOW and (float)bestDist1 < mfNNratio * (float)bestDist2, both initialised to 256, strict '<' updates.
int orbref_db_keyframe_counts(const uint8_t *queries, int nq, const uint8_t *db, int64_t ndb, int kf_rows, int th_low, float ratio,
                              int32_t *counts, int threads) {
    if (kf_rows < 1 || ndb % kf_rows) return ORB_ERR_INVALID;
    const int64_t nkf = ndb / kf_rows;
    auto work = [&](int64_t k0, int64_t k1) {
        for (int64_t k = k0; k < k1; k++) {
            int c = 0;
            for (int q = 0; q < nq; q++) {
                const uint8_t *dq = queries + (size_t)q * 32;
                int bestDist1 = 256, bestDist2 = 256;
                for (int64_t r = k * kf_rows; r < (k + 1) * kf_rows; r++) {
                    const int dist = descriptor_distance_popcnt(dq, db + (size_t)r * 32);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; }
                    else if (dist < bestDist2) { bestDist2 = dist; }
                }
                if (bestDist1 <= th_low && static_cast<float>(bestDist1) < ratio * static_cast<float>(bestDist2)) c++;
            }
            counts[k] = c;
        }
    };
    if (threads <= 1) { work(0, nkf); return ORB_OK; }
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++) th.emplace_back(work, nkf * t / threads, nkf * (t + 1) / threads);
    for (auto &t : th) t.join();
    return ORB_OK;
}

// ref ORBmatcher::ComputeThreeMaxima, ORBmatcher.cc:1604-1645, on bin counts
void orbref_three_maxima(const int *sizes, int L, int *ind1, int *ind2, int *ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    *ind1 = *ind2 = *ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = sizes[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; *ind3 = *ind2; *ind2 = *ind1; *ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; *ind3 = *ind2; *ind2 = i; }
        else if (s > max3) { max3 = s; *ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { *ind2 = -1; *ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { *ind3 = -1; }
}

// rotation bin of ORBmatcher.cc:238-243 (factor = 1/HISTO_LENGTH, the reference's quirk)
int orbref_rot_bin(float angle_a, float angle_b) {
    const int HISTO_LENGTH = 30;
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = angle_a - angle_b;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

// keep[i] = pair i falls in one of the three dominant rotation bins
int orbref_rot_hist_filter(const float *angle_a, const float *angle_b, int n, uint8_t *keep) {
    int sizes[30] = {0};
    std::vector<int> bin(n);
    for (int i = 0; i < n; i++) { bin[i] = orbref_rot_bin(angle_a[i], angle_b[i]); sizes[bin[i]]++; }
    int i1, i2, i3;
    orbref_three_maxima(sizes, 30, &i1, &i2, &i3);
    for (int i = 0; i < n; i++) keep[i] = (bin[i] == i1 || bin[i] == i2 || bin[i] == i3) ? 1 : 0;
    return ORB_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Flat-array Frame (the members the matchers read) and the reference matchers on it.
// ---------------------------------------------------------------------------------------------
#define FRAME_GRID_ROWS 48   // ref Frame.h:39
#define FRAME_GRID_COLS 64   // ref Frame.h:40
static const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   // ref ORBmatcher.cc:37-39

struct orbref_frame {
    int N = 0;
    std::vector<orb_keypoint> mvKeysUn;
    std::vector<uint8_t> mDescriptors;
    std::vector<float> mvuRight;
    float mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<float> mvScaleFactors;
    std::vector<size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];

    // ref Frame::PosInGrid, Frame.cc:387-397
    bool PosInGrid(const orb_keypoint &kp, int &posX, int &posY) const {
        posX = (int)round((kp.x - mnMinX) * mfGridElementWidthInv);
        posY = (int)round((kp.y - mnMinY) * mfGridElementHeightInv);
        if (posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS) return false;
        return true;
    }
    // ref Frame::GetFeaturesInArea, Frame.cc:332-385.  strict = the in-loop level filter of the KeyFrame matchers
    std::vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r, const int minLevel = -1,
                                          const int maxLevel = -1, bool strict = false) const {
        std::vector<size_t> vIndices;
        const int nMinCellX = std::max(0, (int)floor((x - mnMinX - r) * mfGridElementWidthInv));
        if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
        const int nMaxCellX = std::min((int)FRAME_GRID_COLS - 1, (int)ceil((x - mnMinX + r) * mfGridElementWidthInv));
        if (nMaxCellX < 0) return vIndices;
        const int nMinCellY = std::max(0, (int)floor((y - mnMinY - r) * mfGridElementHeightInv));
        if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
        const int nMaxCellY = std::min((int)FRAME_GRID_ROWS - 1, (int)ceil((y - mnMinY + r) * mfGridElementHeightInv));
        if (nMaxCellY < 0) return vIndices;
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<size_t> &vCell = mGrid[ix][iy];
                for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
                    const orb_keypoint &kpUn = mvKeysUn[vCell[j]];
                    if (strict) {
                        if (kpUn.octave < minLevel || kpUn.octave > maxLevel) continue;
                    } else if (bCheckLevels) {
                        if (kpUn.octave < minLevel) continue;
                        if (maxLevel >= 0)
                            if (kpUn.octave > maxLevel) continue;
                    }
                    const float distx = kpUn.x - x;
                    const float disty = kpUn.y - y;
                    if (fabs(distx) < r && fabs(disty) < r) vIndices.push_back(vCell[j]);
                }
            }
        }
        return vIndices;
    }
    const uint8_t *desc(size_t i) const { return &mDescriptors[i * 32]; }
};

static void rot_filter(std::vector<int> *rotHist, int &ind1, int &ind2, int &ind3) {
    int sizes[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
    orbref_three_maxima(sizes, HISTO_LENGTH, &ind1, &ind2, &ind3);
}

extern "C" {

int orbref_frame_create(const orb_frame_desc *d, orbref_frame **out) {
    orbref_frame *f = new orbref_frame();
    f->N = d->n;
    f->mvKeysUn.assign(d->keys_un, d->keys_un + d->n);
    f->mDescriptors.assign(d->descriptors, d->descriptors + (size_t)d->n * 32);
    if (d->u_right) f->mvuRight.assign(d->u_right, d->u_right + d->n);
    else f->mvuRight.assign(d->n, -1.0f);
    f->mnMinX = d->min_x; f->mnMaxX = d->max_x; f->mnMinY = d->min_y; f->mnMaxY = d->max_y;
    // ref Frame.cc:217-218
    f->mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(f->mnMaxX - f->mnMinX);
    f->mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(f->mnMaxY - f->mnMinY);
    f->mvScaleFactors.assign(d->scale_factors, d->scale_factors + d->nlevels);
    for (int i = 0; i < f->N; i++) {   // ref Frame::AssignFeaturesToGrid, Frame.cc:235-250
        int gx, gy;
        if (f->PosInGrid(f->mvKeysUn[i], gx, gy)) f->mGrid[gx][gy].push_back(i);
    }
    if (d->keyframe_view) {   // KeyFrame keeps `int` copies of the bounds (KeyFrame.h:190-193) next to the Frame's grid and cell scales
        f->mnMinX = (float)(int)f->mnMinX; f->mnMaxX = (float)(int)f->mnMaxX; f->mnMinY = (float)(int)f->mnMinY; f->mnMaxY = (float)(int)f->mnMaxY;
    }
    *out = f;
    return ORB_OK;
}
void orbref_frame_destroy(orbref_frame *f) { delete f; }

int orbref_frame_features_in_area(orbref_frame *f, float x, float y, float r, int min_level, int max_level, int32_t *indices,
                                  int capacity, int *n_out) {
    std::vector<size_t> v = f->GetFeaturesInArea(x, y, r, min_level, max_level);
    *n_out = (int)v.size();
    if ((int)v.size() > capacity) return ORB_ERR_CAPACITY;
    for (size_t i = 0; i < v.size(); i++) indices[i] = (int32_t)v[i];
    return ORB_OK;
}

// ref ORBmatcher::SearchForInitialization, ORBmatcher.cc:405-523
int orbref_search_for_initialization(orbref_frame *F1, orbref_frame *F2, float *vbPrevMatched, int32_t *vnMatches12,
                                     int windowSize, float mfNNratio, int mbCheckOrientation, int *nmatches_out) {
    int nmatches = 0;
    const int n1 = F1->N, n2 = F2->N;
    for (int i = 0; i < n1; i++) vnMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> vMatchedDistance(n2, INT_MAX);
    std::vector<int> vnMatches21(n2, -1);
    for (size_t i1 = 0, iend1 = n1; i1 < iend1; i1++) {
        orb_keypoint kp1 = F1->mvKeysUn[i1];
        int level1 = kp1.octave;
        if (level1 > 0) continue;
        std::vector<size_t> vIndices2 = F2->GetFeaturesInArea(vbPrevMatched[2 * i1], vbPrevMatched[2 * i1 + 1], windowSize, level1, level1);
        if (vIndices2.empty()) continue;
        const uint8_t *d1 = F1->desc(i1);
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (std::vector<size_t>::iterator vit = vIndices2.begin(); vit != vIndices2.end(); vit++) {
            size_t i2 = *vit;
            int dist = descriptor_distance(d1, F2->desc(i2));
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = (int)i2; }
            else if (dist < bestDist2) { bestDist2 = dist; }
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * mfNNratio) {
                if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = (int)i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = F1->mvKeysUn[i1].angle - F2->mvKeysUn[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back((int)i1);
                }
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        rot_filter(rotHist, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                int idx1 = rotHist[i][j];
                if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
            }
        }
    }
    for (size_t i1 = 0, iend1 = n1; i1 < iend1; i1++)
        if (vnMatches12[i1] >= 0) {
            vbPrevMatched[2 * i1] = F2->mvKeysUn[vnMatches12[i1]].x;
            vbPrevMatched[2 * i1 + 1] = F2->mvKeysUn[vnMatches12[i1]].y;
        }
    *nmatches_out = nmatches;
    return ORB_OK;
}

// The shared projection-matcher loop on flat queries; see orb_search_windowed in include/orbgpu.h.
// select 0: ORBmatcher.cc:1400-1445 / 1538-1574 / 370-398 (best only);
// select 1: ORBmatcher.cc:76-125 (best/second with octaves and the ratio rule).
int orbref_search_windowed(orbref_frame *F, const orb_window_queries *Q, const uint8_t *claimed_in, int select, int th_dist,
                           float mfNNratio, int mbCheckOrientation, int32_t *assigned, int *nmatches_out) {
    int nmatches = 0;
    std::vector<int> holder(F->N, -1);          // F.mvpMapPoints[i] -> query index, -2 = pre-existing blocking point
    std::vector<uint8_t> blocking(F->N, 0);     // that map point has Observations()>0
    for (int i = 0; i < F->N; i++)
        if (claimed_in && claimed_in[i]) { holder[i] = -2; blocking[i] = 1; }
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    for (int q = 0; q < Q->n; q++) {
        if (Q->valid && !Q->valid[q]) continue;
        const float radius = Q->radius[q];
        std::vector<size_t> vIndices = F->GetFeaturesInArea(Q->x[q], Q->y[q], radius, Q->min_level[q], Q->max_level[q], Q->strict_levels != 0);
        if (vIndices.empty()) continue;
        const uint8_t *dMP = Q->descriptors + (size_t)q * 32;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (std::vector<size_t>::const_iterator vit = vIndices.begin(), vend = vIndices.end(); vit != vend; vit++) {
            const size_t idx = *vit;
            if (holder[idx] != -1)
                if (blocking[idx]) continue;
            if (Q->stereo_ref && F->mvuRight[idx] > 0) {
                const float er = fabs(Q->stereo_ref[q] - F->mvuRight[idx]);
                if (er > Q->stereo_tol[q]) continue;
            }
            const int dist = descriptor_distance(dMP, F->desc(idx));
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F->mvKeysUn[idx].octave; bestIdx = (int)idx;
            } else if (select == 1 && dist < bestDist2) {
                bestLevel2 = F->mvKeysUn[idx].octave; bestDist2 = dist;
            }
        }
        if (bestDist <= th_dist) {
            if (select == 1 && bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
            holder[bestIdx] = q;
            blocking[bestIdx] = Q->claims ? (Q->claims[q] != 0) : 1;
            nmatches++;
            if (mbCheckOrientation) {
                float rot = Q->angle[q] - F->mvKeysUn[bestIdx].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx);
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        rot_filter(rotHist, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i != ind1 && i != ind2 && i != ind3) {
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    holder[rotHist[i][j]] = -3;   // the reference writes NULL here
                    nmatches--;
                }
            }
        }
    }
    for (int i = 0; i < F->N; i++) assigned[i] = holder[i] >= 0 ? holder[i] : (holder[i] == -3 ? -2 : -1);
    *nmatches_out = nmatches;
    return ORB_OK;
}

// ref ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th), ORBmatcher.cc:45-129
int orbref_search_by_projection_points(orbref_frame *F, const orb_track_points *P, const uint8_t *claimed, float th, float nnratio,
                                       int32_t *assigned, int *nmatches) {
    const int n = P->n;
    std::vector<uint8_t> valid(n);
    std::vector<float> radius(n), tol(n);
    std::vector<int32_t> minl(n), maxl(n);
    const bool bFactor = th != 1.0;
    for (int i = 0; i < n; i++) {
        valid[i] = P->in_view[i];
        const int nPredictedLevel = P->level[i];
        float r = P->view_cos[i] > 0.998 ? 2.5f : 4.0f;   // RadiusByViewingCos, ref :131-137
        if (bFactor) r *= th;
        const int lv = std::min(std::max(nPredictedLevel, 0), (int)F->mvScaleFactors.size() - 1);
        radius[i] = r * F->mvScaleFactors[lv];
        tol[i] = r * F->mvScaleFactors[lv];
        minl[i] = nPredictedLevel - 1; maxl[i] = nPredictedLevel;
    }
    orb_window_queries Q;
    memset(&Q, 0, sizeof(Q));
    Q.n = n; Q.valid = valid.data(); Q.x = P->proj_x; Q.y = P->proj_y; Q.radius = radius.data();
    Q.min_level = minl.data(); Q.max_level = maxl.data(); Q.strict_levels = 0;
    Q.stereo_ref = P->proj_xr; Q.stereo_tol = tol.data(); Q.descriptors = P->descriptors; Q.angle = nullptr; Q.claims = P->obs_positive;
    return orbref_search_windowed(F, &Q, claimed, 1, TH_HIGH, nnratio, 0, assigned, nmatches);
}

// ref ORBmatcher::SearchByProjection(Frame &Cur, const Frame &Last, th, bMono), ORBmatcher.cc:1331-1473
int orbref_search_by_projection_last(orbref_frame *Cur, const orb_last_points *P, const uint8_t *claimed, float th, int forward,
                                     int backward, float mbf, int check_orientation, int32_t *assigned, int *nmatches) {
    const int n = P->n;
    std::vector<float> radius(n), sref(n);
    std::vector<int32_t> minl(n), maxl(n);
    for (int i = 0; i < n; i++) {
        const int nLastOctave = P->octave[i];
        radius[i] = th * Cur->mvScaleFactors[nLastOctave];
        if (forward) { minl[i] = nLastOctave; maxl[i] = -1; }
        else if (backward) { minl[i] = 0; maxl[i] = nLastOctave; }
        else { minl[i] = nLastOctave - 1; maxl[i] = nLastOctave + 1; }
        sref[i] = P->u[i] - mbf * P->invz[i];   // ur, ref :1412
    }
    orb_window_queries Q;
    memset(&Q, 0, sizeof(Q));
    Q.n = n; Q.valid = P->valid; Q.x = P->u; Q.y = P->v; Q.radius = radius.data();
    Q.min_level = minl.data(); Q.max_level = maxl.data(); Q.strict_levels = 0;
    Q.stereo_ref = sref.data(); Q.stereo_tol = radius.data(); Q.descriptors = P->descriptors; Q.angle = P->angle; Q.claims = P->obs_positive;
    return orbref_search_windowed(Cur, &Q, claimed, 0, TH_HIGH, 0.f, check_orientation, assigned, nmatches);
}

// ref ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...), ORBmatcher.cc:159-288 (and KF-KF :525-658)
int orbref_search_by_bow(orbref_frame *KF, orbref_frame *F, int ngroups, const int32_t *kf_off, const int32_t *kf_idx,
                         const int32_t *f_off, const int32_t *f_idx, const uint8_t *kf_valid, const uint8_t *f_valid,
                         int strict_low, float mfNNratio, int mbCheckOrientation, int32_t *matches, int *nmatches_out) {
    for (int i = 0; i < F->N; i++) matches[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    for (int g = 0; g < ngroups; g++) {
        for (int a = kf_off[g]; a < kf_off[g + 1]; a++) {
            const unsigned int realIdxKF = kf_idx[a];
            if (!kf_valid[realIdxKF]) continue;
            const uint8_t *dKF = KF->desc(realIdxKF);
            int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
            for (int b = f_off[g]; b < f_off[g + 1]; b++) {
                const unsigned int realIdxF = f_idx[b];
                if (matches[realIdxF] >= 0) continue;
                if (f_valid && !f_valid[realIdxF]) continue;
                const int dist = descriptor_distance(dKF, F->desc(realIdxF));
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                else if (dist < bestDist2) { bestDist2 = dist; }
            }
            if (strict_low ? bestDist1 < TH_LOW : bestDist1 <= TH_LOW) {
                if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                    matches[bestIdxF] = realIdxKF;
                    if (mbCheckOrientation) {
                        float rot = KF->mvKeysUn[realIdxKF].angle - F->mvKeysUn[bestIdxF].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back(bestIdxF);
                    }
                    nmatches++;
                }
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        rot_filter(rotHist, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                matches[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    *nmatches_out = nmatches;
    return ORB_OK;
}

// The candidate loop shared by Fuse (ORBmatcher.cc:895-951 with the reprojection gate :917-945; :1052-1072 without) and
// SearchBySim3 (:1191-1213, :1271-1293): independent queries, `dist<bestDist`, no claims.  See orb_search_window_best.
int orbref_search_window_best(orbref_frame *F, const orb_window_queries *Q, const orb_reproj_gate *gate, int32_t *best_idx,
                              int32_t *best_dist) {
    for (int q = 0; q < Q->n; q++) {
        best_idx[q] = -1; best_dist[q] = INT_MAX;
        if (Q->valid && !Q->valid[q]) continue;
        const float u = Q->x[q], v = Q->y[q], radius = Q->radius[q];
        const std::vector<size_t> vIndices = F->GetFeaturesInArea(u, v, radius, Q->min_level[q], Q->max_level[q], Q->strict_levels != 0);
        if (vIndices.empty()) continue;
        const uint8_t *dMP = Q->descriptors + (size_t)q * 32;
        int bestDist = INT_MAX, bestIdx = -1;
        for (std::vector<size_t>::const_iterator vit = vIndices.begin(), vend = vIndices.end(); vit != vend; vit++) {
            const size_t idx = *vit;
            const orb_keypoint &kp = F->mvKeysUn[idx];
            const int kpLevel = kp.octave;
            if (Q->stereo_ref && F->mvuRight[idx] > 0) {
                const float er = fabs(Q->stereo_ref[q] - F->mvuRight[idx]);
                if (er > Q->stereo_tol[q]) continue;
            }
            if (gate) {
                if (F->mvuRight[idx] >= 0) {   // ref :917-931
                    const float kpx = kp.x, kpy = kp.y, kpr = F->mvuRight[idx];
                    const float ex = u - kpx, ey = v - kpy, er = gate->ur[q] - kpr;
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * gate->inv_level_sigma2[kpLevel] > gate->chi2_stereo) continue;
                } else {                       // ref :932-945
                    const float kpx = kp.x, kpy = kp.y;
                    const float ex = u - kpx, ey = v - kpy;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * gate->inv_level_sigma2[kpLevel] > gate->chi2_mono) continue;
                }
            }
            const int dist = descriptor_distance(dMP, F->desc(idx));
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        best_idx[q] = bestIdx; best_dist[q] = bestDist;
    }
    return ORB_OK;
}

// ref ORBmatcher::CheckDistEpipolarLine, ORBmatcher.cc:140-157
static bool CheckDistEpipolarLine(const orb_keypoint &kp1, const orb_keypoint &kp2, const float *F12, const float *mvLevelSigma2) {
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * mvLevelSigma2[kp2.octave];
}

// ref ORBmatcher::SearchForTriangulation, ORBmatcher.cc:660-826 (this fork never sets vbMatched2)
int orbref_search_for_triangulation(orbref_frame *KF1, orbref_frame *KF2, int ngroups, const int32_t *off1, const int32_t *idx1v,
                                    const int32_t *off2, const int32_t *idx2v, const uint8_t *usable1, const uint8_t *usable2,
                                    const float *F12, float ex, float ey, const float *level_sigma2, int bOnlyStereo,
                                    int mbCheckOrientation, int32_t *vMatches12, int *nmatches_out) {
    int nmatches = 0;
    std::vector<bool> vbMatched2(KF2->N, false);
    for (int i = 0; i < KF1->N; i++) vMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    for (int g = 0; g < ngroups; g++) {
        for (int i1 = off1[g]; i1 < off1[g + 1]; i1++) {
            const size_t idx1 = idx1v[i1];
            if (!usable1[idx1]) continue;
            const bool bStereo1 = KF1->mvuRight[idx1] >= 0;
            if (bOnlyStereo)
                if (!bStereo1) continue;
            const orb_keypoint &kp1 = KF1->mvKeysUn[idx1];
            const uint8_t *d1 = KF1->desc(idx1);
            int bestDist = TH_LOW, bestIdx2 = -1;
            for (int i2 = off2[g]; i2 < off2[g + 1]; i2++) {
                const size_t idx2 = idx2v[i2];
                if (vbMatched2[idx2] || !usable2[idx2]) continue;
                const bool bStereo2 = KF2->mvuRight[idx2] >= 0;
                if (bOnlyStereo)
                    if (!bStereo2) continue;
                const int dist = descriptor_distance(d1, KF2->desc(idx2));
                if (dist > TH_LOW || dist > bestDist) continue;
                const orb_keypoint &kp2 = KF2->mvKeysUn[idx2];
                if (!bStereo1 && !bStereo2) {
                    const float distex = ex - kp2.x;
                    const float distey = ey - kp2.y;
                    if (distex * distex + distey * distey < 100 * KF2->mvScaleFactors[kp2.octave]) continue;
                }
                if (CheckDistEpipolarLine(kp1, kp2, F12, level_sigma2)) { bestIdx2 = (int)idx2; bestDist = dist; }
            }
            if (bestIdx2 >= 0) {
                const orb_keypoint &kp2 = KF2->mvKeysUn[bestIdx2];
                vMatches12[idx1] = bestIdx2;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = kp1.angle - kp2.angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back((int)idx1);
                }
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        rot_filter(rotHist, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                vMatches12[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    *nmatches_out = nmatches;
    return ORB_OK;
}

}  // extern "C"
