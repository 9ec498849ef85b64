#!/bin/bash
# The round's measurement recipe (run under gpurun on one B200): parity suite, bench lines of every workload, the
# reference arm, the ncu launch list of the default bench command and one `ncu --set full` capture per hot kernel.
# Outputs land in gpurun_out/final_*; the summaries kept for the judge are copied to profiles/ by hand.
O=gpurun_out
(time timeout 900 python -m pytest tests -m gpu -x -q) > $O/final_pytest.log 2>&1; grep -E "passed|failed" $O/final_pytest.log
timeout 900 python bench.py > $O/final_bench_default.json 2> $O/final_bench_default.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/final_bench_reference_arm.json 2> $O/final_bench_reference_arm.err
for a in "1080p-m4 iid" "1080p-m8 motion" "4k-m2 motion" "4k-m2 iid" "4k-m0 motion" "4k-m0 iid"; do
  set -- $a
  timeout 900 python bench.py --workload $1 --kind $2 > $O/final_bench_$1_$2.json 2> $O/final_bench_$1_$2.err
done
for f in $O/final_bench_*.json; do python - $f <<'PY'
import json, sys
for line in open(sys.argv[1]):
    if line.startswith("{"):
        d = json.loads(line)
        r = d.get("roofline") or {}
        print(sys.argv[1], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), r.get("kernel"), r.get("frac") and round(r["frac"], 3),
              "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"], 2), [(e["workload"], round(e.get("value", 0), 1), round(e["roofline"]["frac"], 3)) for e in (d.get("extra_workloads") or []) if "roofline" in e])
PY
done
# launch list of the default bench command (numbers printed under ncu are not bench values)
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $O/final_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extra > $O/final_ncu_launches.log 2>&1
# one full capture per hot kernel and workload
ncu --set full --clock-control none --import-source on -k k_hme -s 2 -c 1 -f -o $O/final_k_hme_1080p python tools/profile_step.py 3 motion > $O/final_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k k_fullsearch -s 2 -c 1 -f -o $O/final_k_fullsearch_1080p_iid python tools/profile_step.py 3 iid > $O/final_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k k_hme -s 1 -c 1 -f -o $O/final_k_hme_4k_m0 python tools/profile_step.py 2 motion 4k-m0 > $O/final_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k k_fullsearch -s 1 -c 1 -f -o $O/final_k_fullsearch_4k_m0_iid python tools/profile_step.py 2 iid 4k-m0 > $O/final_ncu4.log 2>&1
ls -la $O | grep final_
