import json, sys
for l in open(sys.argv[1]):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"{d['tag']:8s} {d['scene']} pool {d['pool']:8d} Mpaths/s {d['mpaths_s']:7.1f} (untimed {d['mpaths_s_untimed']:7.1f}) Mrays/s {d['mrays_s']:7.0f} ext {d['stage'][1]*1e3:6.2f} shade {d['stage'][2]*1e3:6.2f} con {d['stage'][3]*1e3:6.2f} ms iters {d['iters']}")
