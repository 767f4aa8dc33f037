#!/bin/bash
# Round artefacts on the GPU box (run through gpurun): the bench line of both arms, the ncu launch list of the bench
# command, full ncu captures of the GEMM / attention kernels (reports stay in gpurun_out/, summaries go to profiles/
# through tools/ncu_summary.py), the MMDT forward sweep and the byte-kernel HBM table.   usage: round_artifacts.sh r1e
tag=${1:-rX}
o=gpurun_out
mkdir -p $o
timeout 600 python bench.py > $o/${tag}_bench_v2a_b8.json 2> $o/${tag}_bench.err; tail -c 600 $o/${tag}_bench_v2a_b8.json; echo
timeout 600 python bench.py --impl reference > $o/${tag}_bench_reference_arm.json 2> $o/${tag}_bench_ref.err; tail -c 400 $o/${tag}_bench_reference_arm.json; echo
timeout 300 python bench.py --direction a2v --batch 64 --steps 5 --warmup 3 --no-cpu-baseline > $o/${tag}_bench_a2v_b64.json 2>> $o/${tag}_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches_ncu.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $o/ncu_launch.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:gemm -s 8 -c 4 -f -o $o/${tag}_gemm_mvp python tools/kbench.py --once > $o/ncu_a.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:gemm -s 8 -c 4 -f -o $o/${tag}_gemm_m32768 python tools/kbench.py --once --M 32768 > $o/ncu_b.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:attn_tc -s 2 -c 1 -f -o $o/${tag}_attn_mvp python tools/kbench.py --once > $o/ncu_c.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:attn_fa -s 2 -c 1 -f -o $o/${tag}_attn_m32768 python tools/kbench.py --once --M 32768 > $o/ncu_d.log 2>&1
for n in gemm_mvp gemm_m32768 attn_mvp attn_m32768; do python tools/ncu_summary.py $o/${tag}_${n}.ncu-rep $o/${tag}_${n}_ncu_full_summary.csv; done
timeout 200 python tools/stepbench.py 8 audio --kinds > $o/${tag}_stepbench.txt 2>&1
timeout 200 python tools/stepbench.py 64 video --kinds >> $o/${tag}_stepbench.txt 2>&1
timeout 200 python tools/stepbench.py 8 video --cfg10 --kinds >> $o/${tag}_stepbench.txt 2>&1
timeout 200 python tools/kbench.py --M 32768 >> $o/${tag}_stepbench.txt 2>&1
timeout 200 python tools/kbench.py --M 17024 >> $o/${tag}_stepbench.txt 2>&1
timeout 200 python tools/kbench.py --M 7120 >> $o/${tag}_stepbench.txt 2>&1
for v in 0 2 3 4; do echo "== AVD_ATTN_FA=$v" >> $o/${tag}_stepbench.txt; AVD_ATTN_FA=$v timeout 200 python tools/kbench.py --M 32768 2>&1 | grep "^attn" >> $o/${tag}_stepbench.txt; done
timeout 900 python tools/sweep.py --precision f16 --out $o/${tag}_sweep_mmdt_forward.json > $o/sweep.log 2>&1; tail -3 $o/sweep.log
timeout 300 python tools/ebench.py --out $o/${tag}_elementwise_hbm.json > $o/ebench.log 2>&1; tail -3 $o/ebench.log
