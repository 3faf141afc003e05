// Developer tool: compiles the two-view device math (essential_math.cuh) for the HOST so that the
// algorithms can be checked against oracle/essential.py in a container without a GPU.
// Not part of the product library and not used by the tests.
#include "../visual_slam_b200/csrc/essential_math.cuh"
using namespace vslam;
extern "C" int hc_five_point(const double* x1, const double* x2, double* models) {
  double a[5][2], b[5][2], m[10][9];
  for (int i = 0; i < 5; ++i) { a[i][0] = x1[2 * i]; a[i][1] = x1[2 * i + 1]; b[i][0] = x2[2 * i]; b[i][1] = x2[2 * i + 1]; }
  int n = five_point(a, b, m);
  for (int i = 0; i < n; ++i) for (int q = 0; q < 9; ++q) models[9 * i + q] = m[i][q];
  return n;
}
extern "C" void hc_decompose(const double* E, double* R1, double* R2, double* t) { decompose_essential(E, R1, R2, t); }
extern "C" void hc_triangulate(const double* R, const double* t, double x1, double y1, double x2, double y2, double* X) {
  triangulate_dlt(R, t, x1, y1, x2, y2, X);
}
