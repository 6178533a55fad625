// Host tables of the BRISK2 sampling pattern for one scale index (see pattern.cc).
#pragma once
#include <vector>

namespace okc {

struct HostPatternPoint { float x, y, sigma; };
struct HostLongPair { unsigned char i, j; short pad; int wdx, wdy; };

struct HostTables {
  int basicscale;
  float scale_value;
  int border;                              // sizeList_[basicscale]
  std::vector<HostPatternPoint> points;    // [1024][66]
  std::vector<unsigned char> short_pairs;  // [384][2] (i, j)
  std::vector<HostLongPair> long_pairs;    // [856]
  std::vector<float> lut;                  // [31][31] layer 0
  std::vector<float> lut_upper;            // [31][31] layers >= 1
};

void build_host_tables(HostTables* t);

}  // namespace okc
