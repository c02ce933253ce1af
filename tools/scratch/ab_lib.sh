# A/B of prebuilt libraries: tools/scratch/ab_lib.sh lib1.so lib2.so ...
for l in "$@"; do
  cp $l ytk-learn_b200/libytkgbdt.so
  echo "== $l"
  bash tools/ab.sh "X=1" 2>&1 | tail -2
done
