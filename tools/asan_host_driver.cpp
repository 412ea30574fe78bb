// Driver for tools/asan_host.sh: walks the host library (fleets, builders, ODH injections, packers, the copy half of
// Copy*Fields, JSON rendering, culler packing) under AddressSanitizer + UBSan + LeakSanitizer.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include "../include/nbgraft_host.h"
int main() {
  setenv("NBH_THREADS", "4", 1);
  for (int kind : {2, 4}) {
    nbh_fleet* f = nbh_fleet_new(kind, 0x6E62677261667405ull + kind, 6000, 100);
    nbh_sts* a; nbh_svc* b; nbh_vs* c;
    if (nbh_fleet_pack(f, nullptr, 0, 6000, &a, &b, &c)) return 1;
    nbg_sts_sparse_batch sv; nbh_sts_view_sparse(a, &sv);
    nbg_svc_gen_batch gv; nbh_svc_view_gen(b, &gv);
    nbg_vs_gen_batch vv; nbh_vs_view_gen(c, &vv);
    for (uint64_t i = 0; i < 300; ++i) {
      nbh_obj *from, *to, *nb;
      for (int which = 1; which <= 5; which += 2) {
        nbh_fleet_get(f, i, which, &from); nbh_fleet_get(f, i, which + 1, &to);
        if (which == 1) nbh_apply_sts_fields(from, to, 0x1F, 1);
        if (which == 3) nbh_apply_svc_fields(from, to);
        if (which == 5) nbh_apply_vs_fields(from, to, 1);
        nbh_obj_digest(to);
        char* j = nbh_obj_to_json(to); nbh_str_free(j);
        nbh_obj_free(from); nbh_obj_free(to);
      }
      nbh_fleet_get(f, i, 0, &nb);
      nbh_obj* m = nbh_obj_clone(nb);
      nbh_odh_mutate(m, 31);
      nbh_obj* s; char err[128];
      nbh_generate_statefulset(m, nullptr, &s, err, sizeof err); nbh_obj_free(s);
      nbh_generate_service(m, &s, err, sizeof err); nbh_obj_free(s);
      nbh_generate_virtualservice(m, nullptr, &s, err, sizeof err); nbh_obj_free(s);
      nbh_obj_free(m); nbh_obj_free(nb);
    }
    nbh_sts_free(a); nbh_svc_free(b); nbh_vs_free(c);
    {  // round 2: lean packing, packing of held objects, namespace tables, table-driven shards, Feast / proxy mutations
      std::vector<uint64_t> nh(100), nsz(100);
      if (nbh_fleet_ns_table(f, nh.data(), nsz.data()) != 100) return 2;
      std::vector<uint32_t> table(100);
      for (int k = 0; k < 100; ++k) table[k] = (uint32_t)(k % 3);
      const uint64_t cnt = nbh_fleet_shard_indices_by_table(f, table.data(), 1, nullptr, 0);
      std::vector<uint64_t> idx(cnt);
      nbh_fleet_shard_indices_by_table(f, table.data(), 1, idx.data(), cnt);
      if (nbh_fleet_pack_lean(f, idx.data(), 0, cnt, &a, &b, &c)) return 3;
      nbg_sts_sparse_batch lv; nbh_sts_view_sparse(a, &lv);
      if (!nbh_sts_is_lean(a)) return 4;
      nbh_sts_free(a); nbh_svc_free(b); nbh_vs_free(c);
      std::vector<nbh_obj*> objs(500 * 7);
      nbh_fleet_objects(f, nullptr, 100, 500, objs.data());
      if (nbh_pack_objects(objs.data(), 500, 1, &a, &b, &c)) return 5;
      nbh_sts_free(a); nbh_svc_free(b); nbh_vs_free(c);
      {  // the desired StatefulSet tile made from the Notebook's: for the device (add_gen) and on the host (add_notebook)
        nbh_sts* g = nbh_sts_new();
        nbh_sts* h = nbh_sts_new();
        for (int k = 0; k < 500; ++k) {
          if (nbh_sts_add_gen(g, objs[k * 7], objs[k * 7 + 2], nullptr)) return 6;
          if (nbh_sts_add_notebook(h, objs[k * 7], objs[k * 7 + 2], nullptr)) return 7;
        }
        nbg_sts_gen_inputs gi; nbg_sts_gen_consts gc; nbg_sts_batch gb;
        if (nbh_sts_view_gen(g, &gi) != 500 || nbh_sts_view_gen(h, &gi) != 0) return 8;
        if (nbh_sts_gen_consts(nullptr, &gc) || nbh_sts_view(h, &gb) != 500) return 9;
        if (nbh_sts_add_notebook(g, objs[0], objs[2], nullptr) == 0) return 10;   // one kind of `from` side per batch
        nbh_sts_free(g); nbh_sts_free(h);
      }
      for (int k = 0; k < 200; ++k) {
        nbh_obj* m = nbh_obj_clone(objs[k * 7]);
        nbh_obj_set_str(m, "metadata.labels.opendatahub.io/feast-integration", k % 2 ? "true" : "false");
        nbh_odh_mutate_proxy(m, 63, "http://p:1", "http://p:2", ".svc");
        nbh_odh_mutate(m, 32);
        nbh_obj_free(m);
      }
      for (nbh_obj* o : objs) nbh_obj_free(o);
    }
    nbh_cullfleet* cf = nbh_cullfleet_new(7, 5000, 1767225600, 0, 4.0);
    nbh_cull* cb; nbh_cullfleet_pack(cf, nullptr, 0, 5000, &cb);
    nbh_cull_free(cb); nbh_cullfleet_free(cf);
    nbh_fleet_free(f);
  }
  printf("asan driver done\n");
  return 0;
}
