"""print the headline numbers of a bench.py JSON line file"""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        st = d["stages"]
        print(f, "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), "check", (d.get("check") or {}).get("ok"),
              "sync", st["sync"]["ms"], "halo_exchange", st["sync.halo_exchange"]["ms"], "halo_discovery",
              st["sync.halo_discovery"]["ms"], "halos", d["observed"]["halos_global"], d["observed"].get("halo_exchange"),
              "|", d["observed"].get("halo_discovery"))
    except Exception as e:  # noqa: BLE001
        print(f, "unreadable:", e, open(f).read()[-800:])
