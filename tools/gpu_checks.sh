#!/bin/bash
# racecheck of the fused kernel on a small batch, larger-robot timing, heterogeneous timing
OUT=gpurun_out/checks; mkdir -p $OUT
python - > $OUT/sanitizer_case_plain.log 2>&1 <<'PY'
import numpy as np, torch
from gmr_b200 import model_pack, synth
from gmr_b200.engine import BatchedRetargeter, DeviceModel
for src, robot in (("bvh","unitree_g1"),("smplx","galaxea_r1pro")):
    m,t = model_pack.build_tables(src, robot)
    human, heights, names = synth.synth_human_clips(m,t,src,B=40,T=6,seed=5)
    dm = DeviceModel.load(src, robot, names); br = BatchedRetargeter(dm, 40); br.set_heights(heights)
    q = br.run(human); torch.cuda.synchronize(); print(robot, float(q.abs().sum()))
PY
echo "plain rc=$?"
cat > /tmp/san_case.py <<'PY'
import numpy as np, torch
from gmr_b200 import model_pack, synth
from gmr_b200.engine import BatchedRetargeter, DeviceModel
for src, robot in (("bvh","unitree_g1"),("smplx","galaxea_r1pro")):
    m,t = model_pack.build_tables(src, robot)
    human, heights, names = synth.synth_human_clips(m,t,src,B=40,T=6,seed=5)
    dm = DeviceModel.load(src, robot, names); br = BatchedRetargeter(dm, 40); br.set_heights(heights)
    q = br.run(human); torch.cuda.synchronize(); print(robot, float(q.abs().sum()))
PY
timeout 600 compute-sanitizer --tool racecheck --print-limit 20 python /tmp/san_case.py > $OUT/racecheck.log 2>&1; echo "racecheck rc=$?"; tail -5 $OUT/racecheck.log
python tools/gpu_quick.py --src smplx --robot unitree_g1_with_hands --B 8192 --T 300 --reps 2 2>&1 | grep -E "launch|rep 1|finite" | tee $OUT/hands_8192.log
python tools/gpu_quick.py --src smplx --robot galaxea_r1pro --B 4096 --T 300 --reps 2 2>&1 | grep -E "launch|rep 1|finite" | tee $OUT/r1pro_4096.log
