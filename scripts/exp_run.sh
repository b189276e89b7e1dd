# usage: bash scripts/exp_run.sh <script.py> tag1 tag2 ...   (runs the script once per experimental library)
S=$1; shift
for t in "$@"; do echo "== $t"; SSPP_CUDA_LIB=$PWD/sspp_b200/csrc/libexp_$t.so python $S 2>&1 | tail -1; done
