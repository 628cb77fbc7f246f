// CPU ORACLE (test infrastructure, not product code) -- matcher side; see match_oracle.h.
#include "match_oracle.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>

namespace oro {

void FrameView::setBounds(float mnX, float mnY, float mxX, float mxY) {
  minX = mnX; minY = mnY; maxX = mxX; maxY = mxY;
  invW = (float)FRAME_GRID_COLS / (maxX - minX);  // Frame.cc:160
  invH = (float)FRAME_GRID_ROWS / (maxY - minY);
}

// Frame::AssignFeaturesToGrid + PosInGrid, Frame.cc:434-465, 815-825.
void FrameView::assignFeaturesToGrid() {
  for (auto& col : grid) for (auto& c : col) c.clear();
  for (int i = 0; i < N; i++) {
    int posX = (int)std::round((kps[i].x - minX) * invW);
    int posY = (int)std::round((kps[i].y - minY) * invH);
    if (posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS) continue;
    grid[posX][posY].push_back(i);
  }
}

// Frame::GetFeaturesInArea, Frame.cc:744-813.
std::vector<int> FrameView::getFeaturesInArea(float x, float y, float r, int minLevel, int maxLevel) const {
  std::vector<int> vIndices;
  const float factorX = r, factorY = r;
  const int nMinCellX = std::max(0, (int)std::floor((x - minX - factorX) * invW));
  if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
  const int nMaxCellX = std::min((int)FRAME_GRID_COLS - 1, (int)std::ceil((x - minX + factorX) * invW));
  if (nMaxCellX < 0) return vIndices;
  const int nMinCellY = std::max(0, (int)std::floor((y - minY - factorY) * invH));
  if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
  const int nMaxCellY = std::min((int)FRAME_GRID_ROWS - 1, (int)std::ceil((y - minY + factorY) * invH));
  if (nMaxCellY < 0) return vIndices;
  const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
  for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
    for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
      for (int j : grid[ix][iy]) {
        const KeyPoint& kp = kps[j];
        if (bCheckLevels) {
          if (kp.octave < minLevel) continue;
          if (maxLevel >= 0 && kp.octave > maxLevel) continue;
        }
        const float distx = kp.x - x, disty = kp.y - y;
        if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) vIndices.push_back(j);
      }
  return vIndices;
}

// ORBmatcher::ComputeThreeMaxima, ORBmatcher.cc:2416-2458.
void compute_three_maxima(const int* hs, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; i++) {
    const int s = hs[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

static inline bool blocks(int slot, const uint8_t* observed) {
  if (slot == -1 || slot == -3) return false;
  if (slot == -2) return true;
  return observed ? observed[slot] != 0 : true;
}

// ORBmatcher.cc:44-142 (the mbTrackInView / Nleft == -1 branch) + RadiusByViewingCos :216-222.
int search_by_projection_mappoints(FrameView& F, int nMP, const uint8_t* in_view, const float* projX, const float* projY,
                                   const int* level, const float* viewCos, const uint8_t* mpDesc, const uint8_t* observed,
                                   const float* projXR, float th, float nnratio, int* kp_mp) {
  int nmatches = 0;
  const bool bFactor = th != 1.0;
  for (int iMP = 0; iMP < nMP; iMP++) {
    if (!in_view[iMP]) continue;
    const int nPredictedLevel = level[iMP];
    float r = (viewCos[iMP] > 0.998) ? 2.5f : 4.0f;
    if (bFactor) r *= th;
    const std::vector<int> vIndices =
        F.getFeaturesInArea(projX[iMP], projY[iMP], r * F.scaleFactors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel);
    if (vIndices.empty()) continue;
    const uint8_t* MPdescriptor = mpDesc + 32 * (size_t)iMP;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int idx : vIndices) {
      if (blocks(kp_mp[idx], observed)) continue;
      if (F.uRight && F.uRight[idx] > 0) {
        const float er = std::fabs(projXR[iMP] - F.uRight[idx]);
        if (er > r * F.scaleFactors[nPredictedLevel]) continue;
      }
      const int dist = hamming256(MPdescriptor, F.desc + 32 * (size_t)idx);
      if (dist < bestDist) {
        bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F.kps[idx].octave; bestIdx = idx;
      } else if (dist < bestDist2) {
        bestLevel2 = F.kps[idx].octave; bestDist2 = dist;
      }
    }
    if (bestDist <= TH_HIGH) {
      if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
      if (bestLevel != bestLevel2 || bestDist <= nnratio * bestDist2) {
        kp_mp[bestIdx] = iMP;
        nmatches++;
      }
    }
  }
  return nmatches;
}

// ORBmatcher.cc:44-211 on a fisheye-stereo frame (F.Nleft != -1): FL = mvKeys / descriptor rows [0, Nleft) / the left grid, FR =
// mvKeysRight / rows [Nleft, N) / mGridRight; kp_mp = F.mvpMapPoints as indices ([nL + nR], coded as above).  Per map point the
// left-camera pass (:63-141; its failed ratio test `continue`s past the right pass, :121-122), then the right-camera pass (:144-211: no
// th factor on the radius, skipped when mnTrackScaleLevelR == -1).  A match also hands the point to the keypoint's stereo partner
// (mvLeftToRightMatch / mvRightToLeftMatch) and counts it (:131-135, :196-200).
int search_by_projection_mappoints_stereo(FrameView& FL, FrameView& FR, int nMP, const uint8_t* inL, const float* pxL, const float* pyL,
                                          const int* lvL, const float* vcL, const uint8_t* inR, const float* pxR, const float* pyR,
                                          const int* lvR, const float* vcR, const uint8_t* mpDesc, const uint8_t* observed,
                                          const int* l2r, const int* r2l, float th, float nnratio, int* kp_mp) {
  int nmatches = 0;
  const bool bFactor = th != 1.0;
  const int Nleft = FL.N;
  for (int iMP = 0; iMP < nMP; iMP++) {
    if (!inL[iMP] && !inR[iMP]) continue;
    const uint8_t* MPdescriptor = mpDesc + 32 * (size_t)iMP;
    if (inL[iMP]) {
      const int nPredictedLevel = lvL[iMP];
      float r = (vcL[iMP] > 0.998) ? 2.5f : 4.0f;
      if (bFactor) r *= th;
      const std::vector<int> vIndices = FL.getFeaturesInArea(pxL[iMP], pyL[iMP], r * FL.scaleFactors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel);
      if (!vIndices.empty()) {
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
          if (blocks(kp_mp[idx], observed)) continue;
          const int dist = hamming256(MPdescriptor, FL.desc + 32 * (size_t)idx);
          if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = FL.kps[idx].octave; bestIdx = idx; }
          else if (dist < bestDist2) { bestLevel2 = FL.kps[idx].octave; bestDist2 = dist; }
        }
        if (bestDist <= TH_HIGH) {
          if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
          if (bestLevel != bestLevel2 || bestDist <= nnratio * bestDist2) {
            kp_mp[bestIdx] = iMP;
            if (l2r[bestIdx] != -1) { kp_mp[l2r[bestIdx] + Nleft] = iMP; nmatches++; }
            nmatches++;
          }
        }
      }
    }
    if (inR[iMP]) {
      const int nPredictedLevel = lvR[iMP];
      if (nPredictedLevel != -1) {
        const float r = (vcR[iMP] > 0.998) ? 2.5f : 4.0f;
        const std::vector<int> vIndices = FR.getFeaturesInArea(pxR[iMP], pyR[iMP], r * FR.scaleFactors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
          if (blocks(kp_mp[idx + Nleft], observed)) continue;
          const int dist = hamming256(MPdescriptor, FR.desc + 32 * (size_t)idx);
          if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = FR.kps[idx].octave; bestIdx = idx; }
          else if (dist < bestDist2) { bestLevel2 = FR.kps[idx].octave; bestDist2 = dist; }
        }
        if (bestDist <= TH_HIGH) {
          if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
          if (r2l[bestIdx] != -1) { kp_mp[r2l[bestIdx]] = iMP; nmatches++; }
          kp_mp[bestIdx + Nleft] = iMP;
          nmatches++;
        }
      }
    }
  }
  return nmatches;
}

// ORBmatcher.cc:2027-2289 (Frame/Frame), :2291-2413 (Frame/KeyFrame) and :489-720 (KeyFrame/Scw, mode 3), search part.
int search_by_projection_window(FrameView& F, int n, const uint8_t* valid, const float* u, const float* v, const int* lvl,
                                const float* ur, const float* angle, const uint8_t* srcDesc, const uint8_t* observed, float th,
                                int mode, int dist_th, bool any_mp_blocks, bool check_orientation, int* kp_mp) {
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];
  const float factor = 1.0f / HISTO_LENGTH;
  for (int i = 0; i < n; i++) {
    if (!valid[i]) continue;
    const int nLastOctave = lvl[i];
    const float radius = th * F.scaleFactors[nLastOctave];
    std::vector<int> vIndices2;
    if (mode == 1) vIndices2 = F.getFeaturesInArea(u[i], v[i], radius, nLastOctave, -1);
    else if (mode == 2) vIndices2 = F.getFeaturesInArea(u[i], v[i], radius, 0, nLastOctave);
    else if (mode == 3) {  // SearchByProjection(KeyFrame*, Scw, ...), ORBmatcher.cc:557-576 / :672-692: octave in [l-1, l]
      for (int i2 : F.getFeaturesInArea(u[i], v[i], radius, -1, -1))
        if (!(F.kps[i2].octave < nLastOctave - 1 || F.kps[i2].octave > nLastOctave)) vIndices2.push_back(i2);
    }
    else vIndices2 = F.getFeaturesInArea(u[i], v[i], radius, nLastOctave - 1, nLastOctave + 1);
    if (vIndices2.empty()) continue;
    const uint8_t* dMP = srcDesc + 32 * (size_t)i;
    int bestDist = 256, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      if (any_mp_blocks ? kp_mp[i2] != -1 : blocks(kp_mp[i2], observed)) continue;
      if (ur && F.uRight && F.uRight[i2] > 0) {
        const float er = std::fabs(ur[i] - F.uRight[i2]);
        if (er > radius) continue;
      }
      const int dist = hamming256(dMP, F.desc + 32 * (size_t)i2);
      if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
    }
    if (bestDist <= dist_th) {
      kp_mp[bestIdx2] = i;
      nmatches++;
      if (check_orientation) {
        float rot = angle[i] - F.kps[bestIdx2].angle;
        if (rot < 0.0) rot += 360.0f;
        int bin = (int)std::round(rot * factor);
        if (bin == HISTO_LENGTH) bin = 0;
        rotHist[bin].push_back(bestIdx2);
      }
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hs[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) hs[i] = (int)rotHist[i].size();
    compute_three_maxima(hs, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (int j : rotHist[i]) { kp_mp[j] = -1; nmatches--; }
  }
  return nmatches;
}

// ORBmatcher.cc:722-837.
int search_for_initialization(const FrameView& F1, FrameView& F2, float* prevMatched, int windowSize, float nnratio,
                              bool check_orientation, int* vnMatches12) {
  int nmatches = 0;
  for (int i = 0; i < F1.N; i++) vnMatches12[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  const float factor = 1.0f / HISTO_LENGTH;
  std::vector<int> vMatchedDistance(F2.N, INT_MAX), vnMatches21(F2.N, -1);
  for (int i1 = 0; i1 < F1.N; i1++) {
    const int level1 = F1.kps[i1].octave;
    if (level1 > 0) continue;
    std::vector<int> vIndices2 = F2.getFeaturesInArea(prevMatched[2 * i1], prevMatched[2 * i1 + 1], (float)windowSize, level1, level1);
    if (vIndices2.empty()) continue;
    const uint8_t* d1 = F1.desc + 32 * (size_t)i1;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      int dist = hamming256(d1, F2.desc + 32 * (size_t)i2);
      if (vMatchedDistance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) bestDist2 = dist;
    }
    if (bestDist <= TH_LOW) {
      if (bestDist < (float)bestDist2 * nnratio) {
        if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
        vnMatches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (check_orientation) {
          float rot = F1.kps[i1].angle - F2.kps[bestIdx2].angle;
          if (rot < 0.0) rot += 360.0f;
          int bin = (int)std::round(rot * factor);
          if (bin == HISTO_LENGTH) bin = 0;
          rotHist[bin].push_back(i1);
        }
      }
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hs[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) hs[i] = (int)rotHist[i].size();
    compute_three_maxima(hs, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i])
        if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
    }
  }
  for (int i1 = 0; i1 < F1.N; i1++)
    if (vnMatches12[i1] >= 0) {
      prevMatched[2 * i1] = F2.kps[vnMatches12[i1]].x;
      prevMatched[2 * i1 + 1] = F2.kps[vnMatches12[i1]].y;
    }
  return nmatches;
}

// Frame.cc:901-1079.
void compute_stereo_matches(int N, const KeyPoint* kpL, const uint8_t* descL, int Nr, const KeyPoint* kpR, const uint8_t* descR,
                            const std::vector<Image>& pyrL, const std::vector<Image>& pyrR, const float* mvScaleFactors,
                            const float* mvInvScaleFactors, StereoParams sp, float* mvuRight, float* mvDepth, int* bestDistOut,
                            int* bestIdxOut) {
  for (int i = 0; i < N; i++) { mvuRight[i] = -1.0f; mvDepth[i] = -1.0f; if (bestDistOut) bestDistOut[i] = -1; if (bestIdxOut) bestIdxOut[i] = -1; }
  const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
  const int nRows = pyrL[0].h;
  std::vector<std::vector<int>> vRowIndices(nRows);
  for (int iR = 0; iR < Nr; iR++) {
    const float kpY = kpR[iR].y;
    const float r = 2.0f * mvScaleFactors[kpR[iR].octave];
    const int maxr = (int)std::ceil(kpY + r);
    const int minr = (int)std::floor(kpY - r);
    for (int yi = minr; yi <= maxr; yi++)
      if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);  // out-of-range rows are UB in the reference
  }
  const float minZ = sp.mb, minD = 0, maxD = sp.mbf / minZ;
  std::vector<std::pair<int, int>> vDistIdx;
  for (int iL = 0; iL < N; iL++) {
    const int levelL = kpL[iL].octave;
    const float vL = kpL[iL].y, uL = kpL[iL].x;
    const size_t row = (size_t)vL;
    if (row >= (size_t)nRows) continue;  // UB in the reference
    const std::vector<int>& vCandidates = vRowIndices[row];
    if (vCandidates.empty()) continue;
    const float minU = uL - maxD, maxU = uL - minD;
    if (maxU < 0) continue;
    int bestDist = TH_HIGH;
    int bestIdxR = 0;
    const uint8_t* dL = descL + 32 * (size_t)iL;
    for (int iR : vCandidates) {
      if (kpR[iR].octave < levelL - 1 || kpR[iR].octave > levelL + 1) continue;
      const float uR = kpR[iR].x;
      if (uR >= minU && uR <= maxU) {
        const int dist = hamming256(dL, descR + 32 * (size_t)iR);
        if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
      }
    }
    if (bestDistOut) bestDistOut[iL] = bestDist;
    if (bestIdxOut) bestIdxOut[iL] = bestDist < TH_HIGH ? bestIdxR : -1;
    if (bestDist < thOrbDist) {
      const float uR0 = kpR[bestIdxR].x;
      const float scaleFactor = mvInvScaleFactors[levelL];
      const float scaleduL = std::round(uL * scaleFactor);
      const float scaledvL = std::round(vL * scaleFactor);
      const float scaleduR0 = std::round(uR0 * scaleFactor);
      const int w = 5;
      const Image& imL = pyrL[levelL];
      const Image& imR = pyrR[levelL];
      const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w);
      if (y0 < 0 || y0 + 2 * w + 1 > imL.h || xl0 < 0 || xl0 + 2 * w + 1 > imL.w) continue;  // cv::Exception in the reference
      short IL[11][11];
      const int cL = imL.row(y0 + w)[xl0 + w];
      for (int yy = 0; yy < 11; yy++) for (int xx = 0; xx < 11; xx++) IL[yy][xx] = (short)(imL.row(y0 + yy)[xl0 + xx] - cL);
      int bestDistS = INT_MAX, bestincR = 0;
      const int L = 5;
      float vDists[2 * 5 + 1];
      const float iniu = scaleduR0 + L - w;
      const float endu = scaleduR0 + L + w + 1;
      if (iniu < 0 || endu >= imR.w) continue;
      bool oob = false;
      for (int incR = -L; incR <= +L; incR++) {
        const int xr0 = (int)(scaleduR0 + incR - w);
        if (xr0 < 0 || xr0 + 2 * w + 1 > imR.w || y0 + 2 * w + 1 > imR.h) { oob = true; break; }  // cv::Exception in the reference
        const int cR = imR.row(y0 + w)[xr0 + w];
        int sum = 0;
        for (int yy = 0; yy < 11; yy++)
          for (int xx = 0; xx < 11; xx++) sum += std::abs((int)IL[yy][xx] - (int)(short)(imR.row(y0 + yy)[xr0 + xx] - cR));
        float dist = (float)(double)sum;
        if (dist < bestDistS) { bestDistS = (int)dist; bestincR = incR; }
        vDists[L + incR] = dist;
      }
      if (oob) continue;
      if (bestincR == -L || bestincR == L) continue;
      const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
      const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
      if (deltaR < -1 || deltaR > 1) continue;
      float bestuR = mvScaleFactors[levelL] * ((float)scaleduR0 + (float)bestincR + deltaR);
      float disparity = (uL - bestuR);
      if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
        mvDepth[iL] = sp.mbf / disparity;
        mvuRight[iL] = bestuR;
        vDistIdx.push_back(std::pair<int, int>(bestDistS, iL));
      }
    }
  }
  if (vDistIdx.empty()) return;  // the reference reads element 0 of an empty vector here
  std::sort(vDistIdx.begin(), vDistIdx.end());
  const float median = vDistIdx[vDistIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;
  for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
    if (vDistIdx[i].first < thDist) break;
    mvuRight[vDistIdx[i].second] = -1;
    mvDepth[vDistIdx[i].second] = -1;
  }
}

// Frame.cc:1246-1253.
void knn2_ratio(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist, uint8_t* good) {
  knn2_hamming(q, nq, t, nt, idx, dist);
  for (int i = 0; i < nq; i++) good[i] = (idx[2 * i + 1] >= 0) && ((float)dist[2 * i] < (float)dist[2 * i + 1] * 0.7);
}

// ---- SURVEY.md §8(f) rank 1 ----------------------------------------------------------------------------------------------------

// Walks two FeatureVectors the way the reference's while-loop does (ORBmatcher.cc:289-299, 440-451): matching node ids in
// ascending order; lower_bound on a sorted map == advance until >=.
template <class Fn>
static void for_common_nodes(const FeatVec& a, const FeatVec& b, Fn fn) {
  int ia = 0, ib = 0;
  while (ia < a.nnodes && ib < b.nnodes) {
    if (a.nodeIds[ia] == b.nodeIds[ib]) { fn(ia, ib); ia++; ib++; }
    else if (a.nodeIds[ia] < b.nodeIds[ib]) { while (ia < a.nnodes && a.nodeIds[ia] < b.nodeIds[ib]) ia++; }
    else { while (ib < b.nnodes && b.nodeIds[ib] < a.nodeIds[ia]) ib++; }
  }
}

static inline int rot_bin_of(float a1, float a2) {  // ORBmatcher.cc:395-401
  const float factor = 1.0f / HISTO_LENGTH;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * factor);
  if (bin == HISTO_LENGTH) bin = 0;
  return bin;
}

// ORBmatcher.cc:273-475 (KF-Frame, Nleft == -1) and :839-979 (KF-KF).
int search_by_bow(int n1, const KeyPoint* kps1, const uint8_t* desc1, const FeatVec& fv1, const uint8_t* elig1, int n2,
                  const KeyPoint* kps2, const uint8_t* desc2, const FeatVec& fv2, const uint8_t* elig2, int maxDist, float nnratio,
                  bool check_orientation, int* matches12) {
  for (int i = 0; i < n1; i++) matches12[i] = -1;
  std::vector<uint8_t> claimed2(n2, 0);
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  for_common_nodes(fv1, fv2, [&](int a, int b) {
    for (int p1 = fv1.nodeStart[a]; p1 < fv1.nodeStart[a + 1]; p1++) {
      const int idx1 = (int)fv1.indices[p1];
      if (idx1 >= n1 || !elig1[idx1]) continue;
      const uint8_t* d1 = desc1 + 32 * (size_t)idx1;
      int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
      for (int p2 = fv2.nodeStart[b]; p2 < fv2.nodeStart[b + 1]; p2++) {
        const int idx2 = (int)fv2.indices[p2];
        if (idx2 >= n2) continue;
        if (claimed2[idx2] || (elig2 && !elig2[idx2])) continue;
        const int dist = hamming256(d1, desc2 + 32 * (size_t)idx2);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
        else if (dist < bestDist2) bestDist2 = dist;
      }
      if (bestDist1 <= maxDist) {
        if ((float)bestDist1 < nnratio * (float)bestDist2) {
          matches12[idx1] = bestIdx2;
          claimed2[bestIdx2] = 1;
          if (check_orientation) rotHist[rot_bin_of(kps1[idx1].angle, kps2[bestIdx2].angle)].push_back(idx1);
          nmatches++;
        }
      }
    }
  });
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hs[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) hs[i] = (int)rotHist[i].size();
    compute_three_maxima(hs, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (int j : rotHist[i]) { matches12[j] = -1; nmatches--; }
  }
  return nmatches;
}

// ORBmatcher.cc:273-475 when the frame is a fisheye-stereo one (F.Nleft != -1, :338-431): side 2 holds N = nLeft2 + Nright features; within
// a node the left-image candidates go through best / second-best, the right-image ones keep their best; the right match is taken inside
// `bestDist1 <= TH_LOW` whatever the left ratio test said, and its own ratio test is voided by "|| true" (:400-403).  Both claim.
int search_by_bow_stereo(int n1, const KeyPoint* kps1, const uint8_t* desc1, const FeatVec& fv1, const uint8_t* elig1, int n2, int nLeft2,
                         const KeyPoint* kps2, const uint8_t* desc2, const FeatVec& fv2, int maxDist, float nnratio, bool check_orientation,
                         int* matches12L, int* matches12R) {
  for (int i = 0; i < n1; i++) { matches12L[i] = -1; matches12R[i] = -1; }
  std::vector<uint8_t> claimed2(n2, 0);
  std::vector<std::pair<int, int> > rotHist[HISTO_LENGTH];   // (idx1, 0 = left match / 1 = right match)
  int nmatches = 0;
  for_common_nodes(fv1, fv2, [&](int a, int b) {
    for (int p1 = fv1.nodeStart[a]; p1 < fv1.nodeStart[a + 1]; p1++) {
      const int idx1 = (int)fv1.indices[p1];
      if (idx1 >= n1 || !elig1[idx1]) continue;
      const uint8_t* d1 = desc1 + 32 * (size_t)idx1;
      int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256, bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
      for (int p2 = fv2.nodeStart[b]; p2 < fv2.nodeStart[b + 1]; p2++) {
        const int idx2 = (int)fv2.indices[p2];
        if (idx2 >= n2 || claimed2[idx2]) continue;
        const int dist = hamming256(d1, desc2 + 32 * (size_t)idx2);
        if (idx2 < nLeft2 && dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = idx2; }
        else if (idx2 < nLeft2 && dist < bestDist2) bestDist2 = dist;
        if (idx2 >= nLeft2 && dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxFR = idx2; }
        else if (idx2 >= nLeft2 && dist < bestDist2R) bestDist2R = dist;
      }
      if (bestDist1 <= maxDist) {
        if ((float)bestDist1 < nnratio * (float)bestDist2) {
          matches12L[idx1] = bestIdxF; claimed2[bestIdxF] = 1;
          if (check_orientation) rotHist[rot_bin_of(kps1[idx1].angle, kps2[bestIdxF].angle)].push_back(std::make_pair(idx1, 0));
          nmatches++;
        }
        if (bestDist1R <= maxDist) {
          matches12R[idx1] = bestIdxFR; claimed2[bestIdxFR] = 1;
          if (check_orientation) rotHist[rot_bin_of(kps1[idx1].angle, kps2[bestIdxFR].angle)].push_back(std::make_pair(idx1, 1));
          nmatches++;
        }
      }
    }
  });
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hs[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) hs[i] = (int)rotHist[i].size();
    compute_three_maxima(hs, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (const auto& j : rotHist[i]) { (j.second ? matches12R : matches12L)[j.first] = -1; nmatches--; }
  }
  return nmatches;
}

// Pinhole.cpp:150-164.
bool epipolar_constrain_pinhole(const KeyPoint& kp1, const KeyPoint& kp2, const float* F, float unc) {
  const float a = kp1.x * F[0] + kp1.y * F[3] + F[6];
  const float b = kp1.x * F[1] + kp1.y * F[4] + F[7];
  const float c = kp1.x * F[2] + kp1.y * F[5] + F[8];
  const float num = a * kp2.x + b * kp2.y + c;
  const float den = a * a + b * b;
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * unc;
}

// ORBmatcher.cc:981-1222.
int search_for_triangulation(int n1, const KeyPoint* kps1, const uint8_t* desc1, const FeatVec& fv1, const uint8_t* elig1,
                             const uint8_t* stereo1, int n2, const KeyPoint* kps2, const uint8_t* desc2, const FeatVec& fv2,
                             const uint8_t* elig2, const uint8_t* stereo2, const float* F12, float epx, float epy,
                             const float* scaleFactors2, const float* levelSigma2_2, bool coarse, bool check_orientation,
                             int* matches12) {
  for (int i = 0; i < n1; i++) matches12[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  for_common_nodes(fv1, fv2, [&](int a, int b) {
    for (int p1 = fv1.nodeStart[a]; p1 < fv1.nodeStart[a + 1]; p1++) {
      const int idx1 = (int)fv1.indices[p1];
      if (!elig1[idx1]) continue;
      const bool bStereo1 = stereo1 && stereo1[idx1];
      const KeyPoint& kp1 = kps1[idx1];
      const uint8_t* d1 = desc1 + 32 * (size_t)idx1;
      int bestDist = TH_LOW, bestIdx2 = -1;
      for (int p2 = fv2.nodeStart[b]; p2 < fv2.nodeStart[b + 1]; p2++) {
        const int idx2 = (int)fv2.indices[p2];
        if (!elig2[idx2]) continue;   // vbMatched2 is never set by the reference (:1026), so only the static gates remain
        const bool bStereo2 = stereo2 && stereo2[idx2];
        const int dist = hamming256(d1, desc2 + 32 * (size_t)idx2);
        if (dist > TH_LOW || dist > bestDist) continue;
        const KeyPoint& kp2 = kps2[idx2];
        if (!bStereo1 && !bStereo2) {
          const float distex = epx - kp2.x, distey = epy - kp2.y;
          if (distex * distex + distey * distey < 100 * scaleFactors2[kp2.octave]) continue;
        }
        if (epipolar_constrain_pinhole(kp1, kp2, F12, levelSigma2_2[kp2.octave]) || coarse) { bestIdx2 = idx2; bestDist = dist; }
      }
      if (bestIdx2 >= 0) {
        matches12[idx1] = bestIdx2;
        nmatches++;
        if (check_orientation) rotHist[rot_bin_of(kp1.angle, kps2[bestIdx2].angle)].push_back(idx1);
      }
    }
  });
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hs[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) hs[i] = (int)rotHist[i].size();
    compute_three_maxima(hs, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (int j : rotHist[i]) { matches12[j] = -1; nmatches--; }
  }
  return nmatches;
}

// ---- KannalaBrandt8::epipolarConstrain -------------------------------------------------------------------------------------------
// Every cv::Mat expression of KannalaBrandt8.cpp:343-444 is written out with the rounding OpenCV gives it (oracle/ref_shim/cv_prims.cc
// holds the same rules behind the stand-in cv:: header the reference's own file is compiled against; tests/test_oracle_cvmat.py pins
// them to cv2): small gemm = float products summed left to right; Mat::dot / cv::norm = double; s*row - row = one rounding
// (addWeighted); Mat / s = multiply by (float)(1.0 / s); cv::SVD = JacobiSVDImpl_<float>.
void kb8_unproject(const float* cam, float u, float v, int libmMode, float* ray) {
  const float pwx = (u - cam[2]) / cam[0], pwy = (v - cam[3]) / cam[1];
  float scale = 1.f;
  float theta_d = sqrtf(pwx * pwx + pwy * pwy);
  theta_d = fminf(fmaxf((float)(-M_PI / 2.f), theta_d), (float)(M_PI / 2.f));
  if (theta_d > 1e-8) {
    float theta = theta_d;
    for (int j = 0; j < 10; j++) {
      const float theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
      const float k0_theta2 = cam[4] * theta2, k1_theta4 = cam[5] * theta4;
      const float k2_theta6 = cam[6] * theta6, k3_theta8 = cam[7] * theta8;
      const float theta_fix = (theta * (1 + k0_theta2 + k1_theta4 + k2_theta6 + k3_theta8) - theta_d) /
                              (1 + 3 * k0_theta2 + 5 * k1_theta4 + 7 * k2_theta6 + 9 * k3_theta8);
      theta = theta - theta_fix;
      if (fabsf(theta_fix) < 1e-6f) break;   // KannalaBrandt8::precision (include/CameraModels/KannalaBrandt8.h:47-52)
    }
    const float tn = libmMode == LIBM_F32 ? tanf(theta) : (float)std::tan((double)theta);
    scale = tn / theta_d;
  }
  ray[0] = pwx * scale; ray[1] = pwy * scale; ray[2] = 1.f;
}

void kb8_project(const float* cam, const float* X, int libmMode, float* u, float* v) {
  const float x2y2 = X[0] * X[0] + X[1] * X[1];
  const float theta = libmMode == LIBM_F32 ? atan2f(sqrtf(x2y2), X[2]) : (float)std::atan2((double)sqrtf(x2y2), (double)X[2]);
  const float psi = libmMode == LIBM_F32 ? atan2f(X[1], X[0]) : (float)std::atan2((double)X[1], (double)X[0]);
  const float theta2 = theta * theta, theta3 = theta * theta2, theta5 = theta3 * theta2, theta7 = theta5 * theta2, theta9 = theta7 * theta2;
  const float r = theta + cam[4] * theta3 + cam[5] * theta5 + cam[6] * theta7 + cam[7] * theta9;
  *u = (float)((double)(cam[0] * r) * std::cos((double)psi) + (double)cam[2]);   // ::cos(double), see oracle/ref_shim/ref_camera.cc
  *v = (float)((double)(cam[1] * r) * std::sin((double)psi) + (double)cam[3]);
}

void jacobi_svd4(const float* A, float* w, float* vt) {
  const int n = 4;
  float At[4][4], Vt[4][4];
  double W[4];
  for (int i = 0; i < n; i++) {
    double sd = 0;
    for (int k = 0; k < n; k++) {
      const float t = A[k * 4 + i];   // At = A^T: row i of At is column i of A
      At[i][k] = t; Vt[i][k] = i == k ? 1.f : 0.f;
      sd += (double)t * t;
    }
    W[i] = sd;
  }
  const float eps = 2 * 1.1920929e-07f;   // FLT_EPSILON * 2
  for (int iter = 0; iter < 30; iter++) {   // max_iter = max(m, 30)
    bool changed = false;
    for (int i = 0; i < n - 1; i++)
      for (int j = i + 1; j < n; j++) {
        float *Ai = At[i], *Aj = At[j];
        double a = W[i], p = 0, b = W[j];
        for (int k = 0; k < n; k++) p += (double)Ai[k] * Aj[k];
        if (std::fabs(p) <= eps * std::sqrt(a * b)) continue;
        p *= 2;
        const double beta = a - b, gamma = hypot(p, beta);
        float c, s;
        if (beta < 0) {
          const double delta = (gamma - beta) * 0.5;
          s = (float)std::sqrt(delta / gamma);
          c = (float)(p / (gamma * s * 2));
        } else {
          c = (float)std::sqrt((gamma + beta) / (gamma * 2));
          s = (float)(p / (gamma * c * 2));
        }
        a = b = 0;
        for (int k = 0; k < n; k++) {
          const float t0 = c * Ai[k] + s * Aj[k];
          const float t1 = -s * Ai[k] + c * Aj[k];
          Ai[k] = t0; Aj[k] = t1;
          a += (double)t0 * t0; b += (double)t1 * t1;
        }
        W[i] = a; W[j] = b;
        changed = true;
        float *Vi = Vt[i], *Vj = Vt[j];
        for (int k = 0; k < n; k++) {
          const float t0 = c * Vi[k] + s * Vj[k];
          const float t1 = -s * Vi[k] + c * Vj[k];
          Vi[k] = t0; Vj[k] = t1;
        }
      }
    if (!changed) break;
  }
  for (int i = 0; i < n; i++) {
    double sd = 0;
    for (int k = 0; k < n; k++) sd += (double)At[i][k] * At[i][k];
    W[i] = std::sqrt(sd);
  }
  for (int i = 0; i < n - 1; i++) {   // selection sort, descending; rows of Vt follow
    int j = i;
    for (int k = i + 1; k < n; k++)
      if (W[j] < W[k]) j = k;
    if (i != j) {
      std::swap(W[i], W[j]);
      for (int k = 0; k < n; k++) std::swap(Vt[i][k], Vt[j][k]);
    }
  }
  for (int i = 0; i < n; i++) {
    w[i] = (float)W[i];
    for (int k = 0; k < n; k++) vt[i * 4 + k] = Vt[i][k];
  }
}

float kb8_triangulate_matches(const Kb8Pair& G, float x1, float y1, float x2, float y2, float sigmaLevel, float unc, int libmMode,
                              float* p3D) {
  float r1[3], r2[3], r21[3];
  kb8_unproject(G.cam1, x1, y1, libmMode, r1);
  kb8_unproject(G.cam2, x2, y2, libmMode, r2);
  const float* R = G.R12;
  for (int i = 0; i < 3; i++) r21[i] = R[3 * i] * r2[0] + R[3 * i + 1] * r2[1] + R[3 * i + 2] * r2[2];   // R12 * r2
  double dot = 0, n1 = 0, n2 = 0;
  for (int i = 0; i < 3; i++) { dot += (double)r1[i] * (double)r21[i]; n1 += (double)r1[i] * r1[i]; n2 += (double)r21[i] * r21[i]; }
  const float cosParallaxRays = (float)(dot / (std::sqrt(n1) * std::sqrt(n2)));
  if (cosParallaxRays > 0.9998) return -1;
  // Tcw1 = [I | 0], Tcw2 = [R21 | t21] with R21 = R12^T, t21 = -R21 * t12 (gemm with alpha = -1)
  float T1[3][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}}, T2[3][4];
  for (int i = 0; i < 3; i++) {
    for (int k = 0; k < 3; k++) T2[i][k] = R[3 * k + i];
    const float t = R[i] * G.t12[0] + R[3 + i] * G.t12[1] + R[6 + i] * G.t12[2];
    T2[i][3] = (float)((double)t * -1.0);
  }
  // A.row(i) = p.x * Tcw.row(2) - Tcw.row(0|1): addWeighted(row2, p, row01, -1), one rounding per element
  float A[16];
  const float ps[4] = {r1[0], r1[1], r2[0], r2[1]};
  for (int i = 0; i < 4; i++) {
    const float(*T)[4] = i < 2 ? T1 : T2;
    for (int k = 0; k < 4; k++) A[4 * i + k] = (float)((double)T[2][k] * (double)ps[i] + (double)T[i & 1][k] * -1.0);
  }
  float w[4], vt[16];
  jacobi_svd4(A, w, vt);
  const float inv = (float)(1.0 / (double)vt[15]);   // x3D.rowRange(0,3) / x3D.at<float>(3)
  const float x3D[3] = {vt[12] * inv, vt[13] * inv, vt[14] * inv};
  const float z1 = x3D[2];
  if (z1 <= 0) return -1;
  const float z2 = (float)(((double)T2[2][0] * x3D[0] + (double)T2[2][1] * x3D[1] + (double)T2[2][2] * x3D[2]) + (double)T2[2][3]);
  if (z2 <= 0) return -1;
  float u, v;
  kb8_project(G.cam1, x3D, libmMode, &u, &v);
  const float errX1 = u - x1, errY1 = v - y1;
  if ((errX1 * errX1 + errY1 * errY1) > 5.991 * sigmaLevel) return -1;
  float x3D2[3];
  for (int i = 0; i < 3; i++) {   // R21 * x3D + t21: small gemm with C
    const float t = T2[i][0] * x3D[0] + T2[i][1] * x3D[1] + T2[i][2] * x3D[2];
    x3D2[i] = (float)((double)t + (double)T2[i][3]);
  }
  kb8_project(G.cam2, x3D2, libmMode, &u, &v);
  const float errX2 = u - x2, errY2 = v - y2;
  if ((errX2 * errX2 + errY2 * errY2) > 5.991 * unc) return -1;
  if (p3D) { p3D[0] = x3D[0]; p3D[1] = x3D[1]; p3D[2] = x3D[2]; }
  return z1;
}

// ORBmatcher.cc:981-1222 with KannalaBrandt8 cameras (one per keyframe, or the two of a fisheye-stereo rig).
int search_for_triangulation_kb8(int n1, const KeyPoint* kps1, const uint8_t* desc1, const FeatVec& fv1, const uint8_t* elig1,
                                 const uint8_t* stereo1, int nLeft1, int n2, const KeyPoint* kps2, const uint8_t* desc2,
                                 const FeatVec& fv2, const uint8_t* elig2, const uint8_t* stereo2, int nLeft2, const Kb8Pair* G,
                                 float epx, float epy, const float* scaleFactors2, const float* levelSigma2_1,
                                 const float* levelSigma2_2, bool coarse, bool check_orientation, int libmMode, int* matches12) {
  for (int i = 0; i < n1; i++) matches12[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  const bool twoCameras = nLeft1 != -1;   // pKF1->mpCamera2 (and pKF2->mpCamera2) set
  for_common_nodes(fv1, fv2, [&](int a, int b) {
    for (int p1 = fv1.nodeStart[a]; p1 < fv1.nodeStart[a + 1]; p1++) {
      const int idx1 = (int)fv1.indices[p1];
      if (!elig1[idx1]) continue;
      const bool bStereo1 = !twoCameras && stereo1 && stereo1[idx1];
      const KeyPoint& kp1 = kps1[idx1];
      const bool bRight1 = twoCameras && idx1 >= nLeft1;
      const uint8_t* d1 = desc1 + 32 * (size_t)idx1;
      int bestDist = TH_LOW, bestIdx2 = -1;
      for (int p2 = fv2.nodeStart[b]; p2 < fv2.nodeStart[b + 1]; p2++) {
        const int idx2 = (int)fv2.indices[p2];
        if (!elig2[idx2]) continue;
        const bool bStereo2 = nLeft2 == -1 && stereo2 && stereo2[idx2];
        const int dist = hamming256(d1, desc2 + 32 * (size_t)idx2);
        if (dist > TH_LOW || dist > bestDist) continue;
        const KeyPoint& kp2 = kps2[idx2];
        const bool bRight2 = nLeft2 != -1 && idx2 >= nLeft2;
        if (!bStereo1 && !bStereo2 && !twoCameras) {
          const float distex = epx - kp2.x, distey = epy - kp2.y;
          if (distex * distex + distey * distey < 100 * scaleFactors2[kp2.octave]) continue;
        }
        const Kb8Pair& g = twoCameras ? G[2 * (int)bRight1 + (int)bRight2] : G[0];
        const bool ok = kb8_triangulate_matches(g, kp1.x, kp1.y, kp2.x, kp2.y, levelSigma2_1[kp1.octave], levelSigma2_2[kp2.octave],
                                                libmMode, nullptr) > 0.0001f;
        if (ok || coarse) { bestIdx2 = idx2; bestDist = dist; }
      }
      if (bestIdx2 >= 0) {
        matches12[idx1] = bestIdx2;
        nmatches++;
        if (check_orientation) rotHist[rot_bin_of(kp1.angle, kps2[bestIdx2].angle)].push_back(idx1);
      }
    }
  });
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hs[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) hs[i] = (int)rotHist[i].size();
    compute_three_maxima(hs, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (int j : rotHist[i]) { matches12[j] = -1; nmatches--; }
  }
  return nmatches;
}

// ORBmatcher.cc:1524-1590 (Fuse), :1728-1756 (Fuse/Scw), :1866-1895 and :1946-1975 (SearchBySim3).
void search_window_best(const FrameView& F, int n, const uint8_t* valid, const float* u, const float* v, const float* radius,
                        const int* level, const float* ur, const uint8_t* desc, int gate, const float* invLevelSigma2,
                        int* bestIdxOut, int* bestDistOut) {
  for (int i = 0; i < n; i++) {
    bestIdxOut[i] = -1; bestDistOut[i] = 256;
    if (!valid[i]) continue;
    const int nPredictedLevel = level[i];
    const std::vector<int> vIndices = F.getFeaturesInArea(u[i], v[i], radius[i], -1, -1);
    const uint8_t* dMP = desc + 32 * (size_t)i;
    int bestDist = 256, bestIdx = -1;
    for (int idx : vIndices) {
      const KeyPoint& kp = F.kps[idx];
      const int kpLevel = kp.octave;
      if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
      if (gate) {
        if (F.uRight && F.uRight[idx] >= 0) {
          const float ex = u[i] - kp.x, ey = v[i] - kp.y, er = ur[i] - F.uRight[idx];
          const float e2 = ex * ex + ey * ey + er * er;
          if (e2 * invLevelSigma2[kpLevel] > 7.8) continue;
        } else {
          const float ex = u[i] - kp.x, ey = v[i] - kp.y;
          const float e2 = ex * ex + ey * ey;
          if (e2 * invLevelSigma2[kpLevel] > 5.99) continue;
        }
      }
      const int dist = hamming256(dMP, F.desc + 32 * (size_t)idx);
      if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
    bestIdxOut[i] = bestIdx; bestDistOut[i] = bestDist;
  }
}


// ---- SURVEY.md §8(f) rank 4 ----------------------------------------------------------------------------------------------------
// MapPoint.cc:385-428.
void distinctive_descriptors(const uint8_t* desc, const int* start, int npoints, int* bestIdxOut, int* bestMedianOut) {
  std::vector<int> vDists;
  for (int p = 0; p < npoints; p++) {
    const int N = start[p + 1] - start[p];
    const uint8_t* d = desc + 32 * (size_t)start[p];
    int BestMedian = INT_MAX, BestIdx = 0;
    if (N <= 0) { bestIdxOut[p] = -1; bestMedianOut[p] = INT_MAX; continue; }
    for (int i = 0; i < N; i++) {
      vDists.resize(N);
      for (int j = 0; j < N; j++) vDists[j] = (i == j) ? 0 : hamming256(d + 32 * (size_t)i, d + 32 * (size_t)j);
      std::sort(vDists.begin(), vDists.end());
      const int median = vDists[(size_t)(0.5 * (N - 1))];
      if (median < BestMedian) { BestMedian = median; BestIdx = i; }
    }
    bestIdxOut[p] = BestIdx; bestMedianOut[p] = BestMedian;
  }
}

// TemplatedVocabulary.h:1218-1259 (FORB::distance = the same bit-count, FORB.cpp:81-101).
void vocab_transform(const VocabTree& V, const uint8_t* desc, int n, int levelsup, int* wordId, int* nodeId, double* weight) {
  const int nid_level = V.L - levelsup;
  for (int f = 0; f < n; f++) {
    const uint8_t* feature = desc + 32 * (size_t)f;
    int nid = 0;   // "if(nid_level <= 0 && nid != NULL) *nid = 0"; otherwise set on the way down
    int final_id = 0, current_level = 0;
    do {
      ++current_level;
      const int c0 = V.childStart[final_id], c1 = V.childStart[final_id + 1];
      final_id = V.childIds[c0];
      double best_d = hamming256(feature, V.nodeDesc + 32 * (size_t)final_id);
      for (int c = c0 + 1; c < c1; c++) {
        const int id = V.childIds[c];
        const double d = hamming256(feature, V.nodeDesc + 32 * (size_t)id);
        if (d < best_d) { best_d = d; final_id = id; }
      }
      if (current_level == nid_level) nid = final_id;
    } while (V.childStart[final_id + 1] > V.childStart[final_id]);   // !isLeaf()
    wordId[f] = V.nodeWordId[final_id];
    nodeId[f] = nid;
    weight[f] = V.nodeWeight[final_id];
  }
}

int vocab_fold(const int* wordId, const int* nodeId, const double* weight, int n, uint32_t* bowIds, double* bowValues,
               uint32_t* fvNodeIds, int* fvStart, uint32_t* fvIndices, int* nnodesOut) {
  std::map<unsigned, double> v;                       // BowVector
  std::map<unsigned, std::vector<unsigned>> fv;       // FeatureVector
  for (int i = 0; i < n; i++) {
    const double w = weight[i];
    if (w > 0) {  // not stopped
      auto vit = v.lower_bound((unsigned)wordId[i]);  // BowVector::addWeight
      if (vit != v.end() && !(v.key_comp()((unsigned)wordId[i], vit->first))) vit->second += w;
      else v.insert(vit, std::make_pair((unsigned)wordId[i], w));
      fv[(unsigned)nodeId[i]].push_back((unsigned)i); // FeatureVector::addFeature
    }
  }
  double norm = 0.0;                                   // BowVector::normalize(L1)
  for (auto& kv : v) norm += std::fabs(kv.second);
  if (norm > 0.0) for (auto& kv : v) kv.second /= norm;
  int k = 0;
  for (auto& kv : v) { bowIds[k] = kv.first; bowValues[k] = kv.second; k++; }
  int nn = 0, pos = 0;
  fvStart[0] = 0;
  for (auto& kv : fv) {
    fvNodeIds[nn] = kv.first;
    for (unsigned f : kv.second) fvIndices[pos++] = f;
    fvStart[++nn] = pos;
  }
  *nnodesOut = nn;
  return k;
}

}  // namespace oro
