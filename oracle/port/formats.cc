// TEST INFRASTRUCTURE ONLY (oracle): CPU restatement of the descriptor formats of SURVEY 8(f) row f4.
// Parity: the text form goes through the SAME iostream calls as the reference's call sites (the semantics live in
// libstdc++, which is what the reference links), so it is as pinned as a restatement can be; meanValue is a
// direct restatement. Nothing under okvis_ceres / okvis_frontend compiles here (Eigen, OpenCV, glog absent).
#include <cstdint>
#include <cstring>
#include <iomanip>
#include <sstream>
#include <string>
#include <vector>

extern "C" {

// okvis_ceres/src/Component.cpp:434-437: `file << setfill('0') << setw(2) << hex << uint32_t(descriptor[i])` x 48
void port_descriptors_to_hex(const uint8_t* desc, int n, char* hex) {
  for (int k = 0; k < n; ++k) {
    std::stringstream file;
    for (size_t i = 0; i < 48; ++i) file << std::setfill('0') << std::setw(2) << std::hex << uint32_t(desc[48 * k + i]);
    const std::string s = file.str();
    std::memcpy(hex + 96 * (size_t)k, s.data(), 96);
  }
}

// okvis_ceres/src/Component.cpp:228-236: per column `stringstream(str.substr(2*col, 2)) >> hex >> uint32_t byte`,
// stored into a uchar. (`byte` is uninitialised there; since C++11 a failed extraction writes 0.)
void port_descriptors_from_hex(const char* hex, int n, uint8_t* desc) {
  for (int k = 0; k < n; ++k) {
    const std::string descriptorstring(hex + 96 * (size_t)k, 96);
    for (int col = 0; col < 48; ++col) {
      uint32_t byte;
      std::stringstream(descriptorstring.substr(2 * col, 2)) >> std::hex >> byte;
      desc[48 * k + col] = (uint8_t)byte;
    }
  }
}

// okvis_frontend/src/FBrisk.cpp:25-60 for each group of rows
void port_descriptor_mean(const uint8_t* desc, const int32_t* group_start, int n_groups, uint8_t* mean) {
  const size_t L = 48;
  for (int g = 0; g < n_groups; ++g) {
    const int begin = group_start[g], end = group_start[g + 1];
    uint8_t* m = mean + L * g;
    std::memset(m, 0, L);
    const uint64_t s = (uint64_t)(end - begin) / 2;
    std::vector<uint64_t> sum(L * 8, 0);
    for (int r = begin; r < end; ++r)
      for (size_t i = 0; i < L; i++)
        for (size_t b = 0; b < 8; ++b) sum[i * 8 + b] += (desc[L * r + i] & (0x01 << b)) ? 1 : 0;
    for (size_t i = 0; i < L; i++)
      for (int b = 0; b < 8; ++b)
        if (sum[i * 8 + size_t(b)] > s) m[i] |= (0x01 << b);
  }
}

}  // extern "C"
