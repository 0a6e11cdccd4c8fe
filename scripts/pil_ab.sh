cd $GRAFT_REPO_ROOT
P=$PWD/ai-laryngeal-video-based-classifier_b200
for n in default pilw6 pilw10; do
  lib=$P/liblvc_b200_$n.so; [ $n = default ] && lib=$P/liblvc_b200.so
  echo "== $n"
  for s in "1080 1920" "720 1280" "480 640" "256 340"; do LVC_B200_LIB=$lib python scripts/perf_hfraw.py vivit_raw $s 2>&1 | tail -1; done
  LVC_B200_LIB=$lib python scripts/perf_hfraw.py timesformer_raw 1080 1920 2>&1 | tail -1
  LVC_B200_LIB=$lib python scripts/perf_hfraw.py timesformer_raw 720 1280 2>&1 | tail -1
done
python -m pytest tests/test_hfraw_gpu.py -x -q -m gpu 2>&1 | tail -1
python tests/fuzz_hfraw.py 100 17 | tail -1
