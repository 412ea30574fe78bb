"""SHA-256 over every array the host packer produces for two seeded fleets (dense + sparse StatefulSet tiles, Service,
VirtualService, their device-generated views) and for the semantic (restart-block) and flattened (ODH comparator)
encodings of the same objects: a packer change that is meant to be a pure speed-up must leave this line unchanged.
usage: python tools/pack_digest.py"""
import hashlib, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("odh-kubeflow_b200")
host = pkg.host
h = hashlib.sha256()
for kind, n in ((host.FLEET_KUBEFLOW, 20000), (host.FLEET_ODH, 8000)):
    f = host.Fleet(kind, host.SEED_BASE + 77, n, 300)
    sts, svc, vs = f.pack()
    for b in (sts, sts.sparse(), svc, svc.gen(), vs, vs.gen()):
        for name, a in sorted(b.arrays().items()):
            h.update(name.encode()); h.update(a.tobytes())
    rb, gen = host.RestartBatch(), host.GenBatch()
    for i in range(0, n, 16):
        nb, a, b2 = f.get(i, "notebook"), f.get(i, "sts_from"), f.get(i, "sts_to")
        gen.add(a, b2)
        rb.add(nb, f.get((i + 1) % n, "notebook"), nb)
    for bb in (gen, rb):
        for name, a in sorted(bb.arrays().items()):
            h.update(name.encode()); h.update(a.tobytes())
print(h.hexdigest())
