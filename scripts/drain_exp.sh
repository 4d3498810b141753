for cfg in "1 2" "2 2" "2 3" "3 3"; do set -- $cfg
echo "== FWD_DRAIN=$1 BWD_DRAIN=$2"
CMF_FWD_DRAIN=$1 CMF_BWD_DRAIN=$2 timeout 300 python -m pytest tests/test_parity_gpu.py -q -k "mnist" 2>&1 | grep -E "passed|failed|mnist_gij_d64 +(grad|loss)|B=50 +(grad|whole|per-param|G|log)" 
CMF_FWD_DRAIN=$1 CMF_BWD_DRAIN=$2 timeout 300 python bench.py --only step --steps 5 --warmup 3
done
