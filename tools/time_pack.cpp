// Packing rate of the host library: nbh_pack_objects over objects that sit in memory, NBH_THREADS threads.
// g++ -O2 -std=c++17 -Iinclude tools/time_pack.cpp -Lodh-kubeflow_b200 -lnbgraft_host -Wl,-rpath,$PWD/odh-kubeflow_b200 -o build/time_pack
// NBH_THREADS=1 build/time_pack 20000 5
#include "nbgraft_host.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
int main(int argc, char** argv) {
  uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 20000;
  int reps = argc > 2 ? atoi(argv[2]) : 5;
  nbh_fleet* f = nbh_fleet_new(2, 1234, n, n / 50);
  std::vector<nbh_obj*> objs(n * 7);
  if (nbh_fleet_objects(f, nullptr, 0, n, objs.data()) != 0) return 1;
  nbh_sts* s; nbh_svc* v; nbh_vs* w;
  auto t0 = std::chrono::steady_clock::now();
  for (int r = 0; r < reps; ++r) {
    if (nbh_pack_objects(objs.data(), n, 1, &s, &v, &w) != 0) return 1;
    nbh_sts_free(s); nbh_svc_free(v); nbh_vs_free(w);
  }
  double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  printf("pack_objects wall: %.2f us per notebook (x threads)\n", dt / reps / n * 1e6);
  return 0;
}
