python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "logpolar or scalerot" 2>&1 | tail -1
for n in 296 2368; do echo "A: $(python tools/prof_side.py $n | head -1)"; echo "B640/2: $(MOF_B200_LIB=$PWD/ab_variant.so python tools/prof_side.py $n | head -1)"; echo "B768/2: $(MOF_B200_LIB=$PWD/ab_v768.so python tools/prof_side.py $n | head -1)"; done
MOF_B200_LIB=$PWD/ab_v768.so python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "logpolar or scalerot" 2>&1 | tail -1
