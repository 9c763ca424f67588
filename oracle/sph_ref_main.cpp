// oracle/sph_ref_main.cpp -- test infrastructure.  Drives the vendored sphericart CPU implementation
// (/root/reference/submodules/sphericart, pinned 1866da4) to emit golden vectors for the l<=3 basis and its
// normalised gradients.  Compiled by oracle/gen_golden.py in the build container only; output is committed
// as tests/golden/sph_l3.json.  No reference source is copied: the library is compiled where it lies.
#include <cstdio>
#include <vector>
#include "sphericart.hpp"
int main() {
  const int l_max = 3, n = 16, N = 64;
  std::vector<double> xyz(N * 3);
  unsigned long long s = 0x3D65ULL;
  auto rnd = [&]() { s = s * 6364136223846793005ULL + 1442695040888963407ULL; return ((s >> 11) * (1.0 / 9007199254740992.0)) * 2.0 - 1.0; };
  for (auto &v : xyz) v = rnd() * 3.0;
  sphericart::SphericalHarmonics<double> calc(l_max);   // normalised variant, as sphericart::cuda::SphericalHarmonics
  std::vector<double> sph(N * n), dsph(N * 3 * n);
  calc.compute_with_gradients(xyz, sph, dsph);
  printf("{\"l_max\":3,\"xyz\":[");
  for (int i = 0; i < N * 3; ++i) printf("%s%.17g", i ? "," : "", xyz[i]);
  printf("],\"sph\":[");
  for (int i = 0; i < N * n; ++i) printf("%s%.17g", i ? "," : "", sph[i]);
  printf("],\"dsph\":[");
  for (int i = 0; i < N * 3 * n; ++i) printf("%s%.17g", i ? "," : "", dsph[i]);
  printf("]}\n");
  return 0;
}
