for v in turn1 turn3 turn4 nopre; do echo "== variant $v"; ITA_B200_LIB=$PWD/build_tools/$v/libita_b200.so timeout 300 python tools/mode_bench.py --quick 2>&1 | cut -c1-70; done
echo "== this build"; timeout 300 python tools/mode_bench.py --quick 2>&1 | cut -c1-70
echo "== torch_fast 0"; ITA_FUSED_TORCH_FAST=0 timeout 300 python tools/mode_bench.py --quick 2>&1 | cut -c1-70 | sed -n 2p
