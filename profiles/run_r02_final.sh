#!/bin/bash
# round 2, final state: whole GPU suite, smoke(), bench.py (default flags) and the reference arm, and the RASR-side
# check binary's own report (scorers and GpuAligner against the reference classes in process, with its timing lines)
mkdir -p gpurun_out
O=gpurun_out/final_summary.txt
: > $O
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/final_suite.log 2>&1; echo "suite rc=$?" >> $O; tail -4 gpurun_out/final_suite.log >> $O
python -c "import __graft_entry__ as g; g.smoke()" >> $O 2>&1; echo "smoke rc=$?" >> $O
python bench.py > gpurun_out/bench_n1_r02.json 2> gpurun_out/bench_n1_r02.err; echo "bench rc=$?" >> $O
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_n1_r02.json 2> gpurun_out/bench_ref_n1_r02.err; echo "reference arm rc=$?" >> $O
python - >> $O 2>&1 <<'PY'
import subprocess, tempfile, os
from rasr_b200 import synth
ms = synth.make_model(M=4501, K=16, D=45)
x = synth.make_features(ms, 400)
with tempfile.TemporaryDirectory() as tmp:
    ms.save(os.path.join(tmp, "m.gms")); x.tofile(os.path.join(tmp, "x.f32"))
    r = subprocess.run(["oracle/_ref/rasr_gpu_check", f"model={tmp}/m.gms", f"features={tmp}/x.f32", "buffer-size=0",
                        "timing-frames=4000"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
print("rasr_gpu_check rc=%d (4501 x 16 x 45, 400 frames; timing on 4000-frame segments)" % r.returncode)
for line in r.stdout.splitlines():
    if " PASS " in line or " FAIL " in line or " SKIP " in line or line.startswith(("TIMING", "ALIGN-TIMING")):
        print(line)
PY
cat $O
head -c 600 gpurun_out/bench_n1_r02.json
