// Host cost of the three ways to put a (desired, live) StatefulSet pair into a batch, one thread, through the host C ABI:
// generateStatefulSet + nbh_sts_add, nbh_sts_add_notebook (desired tile patched on the host), nbh_sts_add_gen (patched
// on the device).  g++ -O2 -std=c++17 -Iinclude tools/time_host_gen.cpp -Lodh-kubeflow_b200 -lnbgraft_host
//   -Wl,-rpath,$PWD/odh-kubeflow_b200 -o build/time_host_gen && build/time_host_gen 20000
#include "nbgraft_host.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main(int argc, char** argv) {
  uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 20000;
  nbh_fleet* f = nbh_fleet_new(2, 1234, n, n / 50);
  std::vector<nbh_obj*> objs(n * 7);
  if (nbh_fleet_objects(f, nullptr, 0, n, objs.data()) != 0) return 1;
  for (int rep = 0; rep < 3; ++rep) {
    double t0 = now();
    nbh_sts* a = nbh_sts_new();
    for (uint64_t i = 0; i < n; ++i) {
      nbh_obj* ss = nullptr; char err[64];
      if (nbh_generate_statefulset(objs[i * 7], nullptr, &ss, err, sizeof err) != 0) return 2;
      nbh_sts_add(a, ss, objs[i * 7 + 2]);
      nbh_obj_free(ss);
    }
    double t1 = now();
    nbh_sts* b = nbh_sts_new();
    for (uint64_t i = 0; i < n; ++i) if (nbh_sts_add_notebook(b, objs[i * 7], objs[i * 7 + 2], nullptr) != 0) return 3;
    double t2 = now();
    nbh_sts* c = nbh_sts_new();
    for (uint64_t i = 0; i < n; ++i) if (nbh_sts_add_gen(c, objs[i * 7], objs[i * 7 + 2], nullptr) != 0) return 4;
    double t3 = now();
    printf("per notebook, one thread: generate+add %.2f us   add_notebook %.2f us   add_gen %.2f us\n",
           (t1 - t0) / n * 1e6, (t2 - t1) / n * 1e6, (t3 - t2) / n * 1e6);
    nbh_sts_free(a); nbh_sts_free(b); nbh_sts_free(c);
  }
  return 0;
}
