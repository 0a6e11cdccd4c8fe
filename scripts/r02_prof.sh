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
python scripts/clip_knobs.py swin_cfg4 timesformer_cfg3 vivit_cfg2 > $O/knobs_default.txt 2>&1
LVC_B200_LIB=$GRAFT_REPO_ROOT/ai-laryngeal-video-based-classifier_b200/liblvc_b200_g4.so CLIP_KNOBS_JSON='[{}, {"LVC_CLIP_RPS": "4"}, {"LVC_CLIP_RPS": "16"}]' python scripts/clip_knobs.py swin_cfg4 timesformer_cfg3 > $O/knobs_g4.txt 2>&1
LVC_B200_LIB=$GRAFT_REPO_ROOT/ai-laryngeal-video-based-classifier_b200/liblvc_b200_mb2.so CLIP_KNOBS_JSON='[{}, {"LVC_CLIP_SMEM_KB": "112"}, {"LVC_CLIP_SMEM_KB": "112", "LVC_CLIP_RPS": "4"}, {"LVC_CLIP_SMEM_KB": "112", "LVC_CLIP_RPS": "2"}]' python scripts/clip_knobs.py swin_cfg4 timesformer_cfg3 > $O/knobs_mb2.txt 2>&1
ls -la $O
