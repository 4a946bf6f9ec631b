// oracle/textio.cpp — TEST INFRASTRUCTURE (see pcp_oracle.h).
//
// preprocess::txt2pc (src/smoothing.cpp:39-81) with the reference's own statements: getline per line, a
// std::stringstream per line, four `ss >> float` extractions into ONE point object that lives across lines
// (a field the line does not supply keeps the previous line's value; a field that fails to parse becomes 0
// and stops the line), push_back per line.  The try/catch upstream never fires (stream exceptions are off).
// The per-line std::cout print is left out.  The C++ library of this image is the arithmetic (num_get ->
// strtof, correctly rounded): nothing is restated.
#include "pcp_oracle.h"

#include <sstream>
#include <string>

extern "C" int64_t orc_txt2pc(const char* text, int64_t n_bytes, float* out_xyzi, int64_t capacity) {
  std::istringstream txtfile(std::string(text, (size_t)n_bytes));
  float pt[4] = {0.f, 0.f, 0.f, 0.f};  // pcl::PointXYZI pt; (x = y = z = 0, intensity = 0)
  std::string line;
  int64_t count = 0;
  while (getline(txtfile, line)) {
    std::stringstream ss(line);
    ss >> pt[0];
    ss >> pt[1];
    ss >> pt[2];
    ss >> pt[3];
    if (count < capacity)
      for (int a = 0; a < 4; ++a) out_xyzi[4 * count + a] = pt[a];
    ++count;
  }
  return count;
}

// preprocess::xyz2pc (src/smoothing.cpp:83-125), the reference's own statements on this image's C library:
//   while (!xyzfile.eof()) { xyzfile.getline(line, 256); sscanf(line, "%f %f %f %s", &pt.x, &pt.y, &pt.z, &intens);
//                            pt.intensity = intens; origiCloud.push_back(pt); }
// with ONE change that keeps the run defined: the reference scans "%s" into a single `char` (any token overflows it
// by at least the terminating NUL); here the token lands in a line-sized buffer and its first byte is what the
// reference's `intens` would hold.  `intens` and `pt` live across lines as upstream (fields a line does not
// supply keep their previous values; the reference's are uninitialised at first, here 0).  Returns the number of
// points, or -1 when a line has 256 characters or more (getline(line, 256) then sets failbit and the reference's
// loop never ends).
#include <cstdio>
extern "C" int64_t orc_xyz2pc(const char* text, int64_t n_bytes, float* out_xyzi, int64_t capacity) {
  std::istringstream xyzfile(std::string(text, (size_t)n_bytes));
  float pt[4] = {0.f, 0.f, 0.f, 0.f};
  char line[256];
  char token[256];
  char intens = 0;
  int64_t count = 0;
  while (!xyzfile.eof()) {
    xyzfile.getline(line, 256);
    if (xyzfile.fail() && !xyzfile.eof()) return -1;  // more than 255 characters before the newline
    token[0] = intens;
    const int got = std::sscanf(line, "%f %f %f %255s", &pt[0], &pt[1], &pt[2], token);
    if (got == 4) intens = token[0];
    pt[3] = intens;  // char -> float (signed on this platform, as on the reference's)
    if (count < capacity)
      for (int a = 0; a < 4; ++a) out_xyzi[4 * count + a] = pt[a];
    ++count;
  }
  return count;
}
