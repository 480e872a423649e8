for v in v0 v4 v5 v0 v4 v5; do echo "== $v"; GUDA_B200_LIB=$PWD/scratch/variants/lib$v.so GUDA_B200_LIB_PARTIAL=1 python scratch/quick_sgemm.py 2>&1 | grep -E "sgemm (1024|4096|8192)"; done
