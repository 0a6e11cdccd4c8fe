# A/B of alternative builds of the library (profiling aid): bash scripts/ab_libs.sh name1 name2 ...  (liblvc_b200_<name>.so)
cd ${GRAFT_REPO_ROOT:-.}
P=ai-laryngeal-video-based-classifier_b200
for n in "$@"; do
  lib=$PWD/$P/liblvc_b200_$n.so
  [ "$n" = default ] && lib=$PWD/$P/liblvc_b200.so
  echo "== $n"
  LVC_B200_LIB=$lib CLIP_KNOBS_JSON='[{}]' python scripts/clip_knobs.py 2>&1 | tail -3
  [ -n "$AB_FUZZ" ] && LVC_B200_LIB=$lib python tests/fuzz_parity.py $AB_FUZZ 31 2>&1 | tail -1
done
