set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:lvc_ -c 400 --csv --log-file $O/r02_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $O/r02_launches.out 2>&1
ncu --set full --clock-control none --import-source on -k regex:lvc_clip -s 3 -c 1 -f -o $O/r02_clip_headline python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-secondary > $O/ncu_headline.out 2>&1
for c in swin_cfg4 vivit_cfg2 timesformer_cfg3; do
  ncu --set full --clock-control none --import-source on -k regex:lvc_clip -s 3 -c 1 -f -o $O/r02_clip_$c python scripts/perf_clip_cfg.py $c > $O/ncu_$c.out 2>&1
done
for m in timesformer vivit; do
  ncu --set full --clock-control none --import-source on -k regex:gemm2 -s 6 -c 1 -f -o $O/r02_gemm2_$m python scripts/perf_embed.py $m > $O/ncu_gemm_$m.out 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:lvc_letterbox -s 3 -c 1 -f -o $O/r02_letterbox python scripts/perf_letterbox.py > $O/ncu_letterbox.out 2>&1
python scripts/clip_knobs.py swin_cfg4 timesformer_cfg3 vivit_cfg2 > $O/knobs_default.txt 2>&1
ls -la $O
