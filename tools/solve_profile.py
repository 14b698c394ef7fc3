#!/usr/bin/env python
"""Per-tag device/host time of one fea config on the GPU path:
    python tools/solve_profile.py bob.json [override.json ...]
(config names are resolved in oracle/_ref/config)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sanm_b200  # noqa: E402
from sanm_b200 import fea  # noqa: E402

CONFIG_DIR = os.path.join(ROOT, "oracle", "_ref", "config")


def main():
    b = sanm_b200.load()
    if os.environ.get("SANM_PROFILE_PREWARM"):
        # warm the CUDA context and the kernel images with another config, so
        # that the cold run below shows topology-dependent costs only
        fea.run_task(b, os.path.join(CONFIG_DIR, os.environ["SANM_PROFILE_PREWARM"]),
                     [], save=False, log=fea.RunLog(verbose=False))
        from sanm_b200.capi import Binding as _B
        _B.call_times.clear()
    task = os.path.join(CONFIG_DIR, sys.argv[1])
    ovs = [os.path.join(CONFIG_DIR, o) for o in sys.argv[2:]]
    for rep in range(2):  # second run: warm allocator / module load
        b.prof_reset()
        t0 = time.perf_counter()
        res = fea.run_task(b, task, ovs, save=False,
                           log=fea.RunLog(verbose=False))
        wall = time.perf_counter() - t0
        if rep == 0 and os.environ.get("SANM_PROFILE_COLD"):
            print("---- cold run: wall_s", wall)
            print(b.prof_dump())
            from sanm_b200.capi import Binding
            for k, (t, n) in sorted(Binding.call_times.items(),
                                    key=lambda kv: -kv[1][0])[:12]:
                print("  call %-34s %9.2f ms  n=%d" % (k, t * 1e3, n))
            Binding.call_times.clear()
    st = res[1]
    print("wall_s", wall)
    print("stats", st if not isinstance(st, list) else
          [{k: s[k] for k in ("iter_tot", "time")} for s in st])
    print(b.prof_dump())


if __name__ == "__main__":
    main()
