TAG=${1:-x3}
mkdir -p gpurun_out
TS2D_TC_DBG=2 TS2D_KSPLIT=9 timeout 300 python tools/ncu_forward.py --forwards 2 > /dev/null 2> gpurun_out/${TAG}_dbg_s9_warm.txt
TS2D_TC_DBG=2 TS2D_KSPLIT=0 timeout 300 python tools/ncu_forward.py --forwards 2 > /dev/null 2> gpurun_out/${TAG}_dbg_s0_warm.txt
TS2D_TC_DBG=1 TS2D_KSPLIT=9 TS2D_L2PROMO_W=0 TS2D_L2PROMO=0 TS2D_L2PROMO_SUB=0 timeout 300 python tools/ncu_forward.py --forwards 2 > /dev/null 2> gpurun_out/${TAG}_dbg_s9_promo0.txt
TS2D_TC_DBG=1 TS2D_KSPLIT=9 TS2D_KSPLIT_BN=64 timeout 300 python tools/ncu_forward.py --forwards 2 > /dev/null 2> gpurun_out/${TAG}_dbg_s9_bn64.txt
TS2D_TC_DBG=1 TS2D_KSPLIT=3 TS2D_KSPLIT_BN=64 timeout 300 python tools/ncu_forward.py --forwards 2 > /dev/null 2> gpurun_out/${TAG}_dbg_s3_bn64.txt
for f in s9_warm s0_warm s9_promo0 s9_bn64 s3_bn64; do echo "== $f"; grep tc_dbg gpurun_out/${TAG}_dbg_$f.txt | tail -31 | grep -E "enc5.1|enc6.0|enc7.0|dec0.0" | cut -c1-160; done
