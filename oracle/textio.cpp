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
