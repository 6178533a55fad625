// Host-side construction of the BRISK2 look-up tables, once per context.
//
// Mirrors brisk::BriskDescriptorExtractor::InitFromStream (brisk/src/brisk-descriptor-extractor.cc:189-302)
// expression by expression, with the host libm the reference itself uses (cos, sin, pow, log, sqrt), but only for
// the single scale index okvis ever uses (scaleInvariant = false -> "basicscale", :586-600), and the layer-0
// uniformity LUT of ScaleSpaceLayer::Create (brisk/include/brisk/internal/scale-space-layer-inl.h:98-107).
// This is init-time table generation, not a compute fallback: the tables are uploaded and only read by kernels.
#include <algorithm>
#include <cmath>
#include <vector>

#include "pattern.h"
#include "brisk_pattern_data.inc"

namespace okc {

void build_host_tables(HostTables* t) {
  const float basicSize = 12.0f;       // basicSize_
  const unsigned int scales = 64;      // scales_
  const float scalerange = 30.0f;      // scalerange_
  const unsigned int n_rot = 1024;     // n_rot_
  const unsigned int points = OKC_PATTERN_POINTS;

  // :586-600 (float statics, unqualified log -> double log of the promoted float)
  const float log2f_ = 0.693147180559945;
  const float lb_scalerange = log((double)scalerange) / (log2f_);
  const float basicSize06 = basicSize * 0.6;
  const int basicscale =
      std::max(static_cast<int>(scales / lb_scalerange * (log(1.45 * basicSize / (basicSize06)) / log2f_) + 0.5), 0);
  t->basicscale = basicscale;

  // :204-208
  const float lb_scale = log((double)scalerange) / log(2.0);
  const float lb_scale_step = lb_scale / (scales);
  const float sigma_scale = 1.3;
  const unsigned int scale = static_cast<unsigned int>(basicscale);
  const float scaleVal = pow(2.0, static_cast<double>(scale * lb_scale_step));
  t->scale_value = scaleVal;

  t->points.resize(static_cast<size_t>(n_rot) * points);
  unsigned int sizeMax = 0;
  for (size_t rot = 0; rot < n_rot; ++rot) {
    for (unsigned int i = 0; i < points; i++) {
      const float ux = kPatternXYS[i][0], uy = kPatternXYS[i][1], us = kPatternXYS[i][2];
      const double theta = static_cast<double>(rot) * 2 * M_PI / static_cast<double>(n_rot);
      HostPatternPoint p;
      p.x = scaleVal * (ux * cos(theta) - uy * sin(theta));
      p.y = scaleVal * (ux * sin(theta) + uy * cos(theta));
      p.sigma = sigma_scale * scaleVal * us;
      const unsigned int size = ceil(((sqrt((double)(p.x * p.x + p.y * p.y))) + p.sigma)) + 1;
      if (sizeMax < size) sizeMax = size;
      t->points[rot * points + i] = p;
    }
  }
  t->border = static_cast<int>(sizeMax);

  t->short_pairs.resize(OKC_SHORT_PAIRS * 2);
  for (int p = 0; p < OKC_SHORT_PAIRS; ++p) {
    t->short_pairs[2 * p] = kShortPairs[p][0];
    t->short_pairs[2 * p + 1] = kShortPairs[p][1];
  }
  t->long_pairs.resize(OKC_LONG_PAIRS);
  for (int p = 0; p < OKC_LONG_PAIRS; ++p) {
    const unsigned i = kLongPairs[p][0], j = kLongPairs[p][1];
    float dx = (kPatternXYS[j][0] - kPatternXYS[i][0]);
    float dy = (kPatternXYS[j][1] - kPatternXYS[i][1]);
    float norm_sq = dx * dx + dy * dy;
    HostLongPair lp;
    lp.i = static_cast<unsigned char>(i);
    lp.j = static_cast<unsigned char>(j);
    lp.pad = 0;
    lp.wdx = static_cast<int>((dx / (norm_sq)) * 2048.0 + 0.5);
    lp.wdy = static_cast<int>((dy / (norm_sq)) * 2048.0 + 0.5);
    t->long_pairs[p] = lp;
  }

  // layer-0 uniformity LUT: max(1 - d^(1/8 of squared distance) / 15^(1/4), 0), double then float
  t->lut.resize(31 * 31);
  for (int x = 0; x < 31; ++x)
    for (int y = 0; y < 31; ++y)
      t->lut[y * 31 + x] = std::max(
          1 - sqrt(sqrt(sqrt(static_cast<double>((15 - x) * (15 - x) + (15 - y) * (15 - y))))) /
                  sqrt(sqrt(sqrt(static_cast<double>(15 * 15)))),
          0.0);
  // upper layers (scale-space-layer-inl.h:185-193): max(1 - d^2 / 15^2, 0)
  t->lut_upper.resize(31 * 31);
  for (int x = 0; x < 31; ++x)
    for (int y = 0; y < 31; ++y)
      t->lut_upper[y * 31 + x] = static_cast<float>(std::max(
          1 - static_cast<double>((15 - x) * (15 - x) + (15 - y) * (15 - y)) / static_cast<double>(15 * 15), 0.0));
}

}  // namespace okc
