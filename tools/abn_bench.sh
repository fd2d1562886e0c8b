#!/bin/bash
# Same-box comparison of several builds of libfa_b200 (boxes differ by a few percent, so variants must share a call):
#   ROUNDS=2 tools/abn_bench.sh libfa_b200.so libfa_b200_X.so libfa_b200_Y.so ...
# alternates the libraries ROUNDS times over tools/prefill_bench.py and prints per-config means relative to the first.
R=${ROUNDS:-2}
D=vllm-flash-attention_b200/vllm_flash_attn
for r in $(seq 1 $R); do
  for lib in "$@"; do
    FA_B200_LIB=$PWD/$D/$lib timeout 200 python tools/prefill_bench.py --iters 30 2>/dev/null | grep '"config"' | sed "s/^/$lib /"
  done
done | python -c "
import sys, json, collections
d = collections.defaultdict(list); libs = []
for ln in sys.stdin:
    v, js = ln.split(' ', 1); o = json.loads(js); d[(o['config'], v)].append(o['TFLOPs'])
    if v not in libs: libs.append(v)
cfgs = sorted({c for c, _ in d})
print(' ' * 46 + ''.join(f'{l[10:-3] or \"base\":>14s}' for l in libs))
for c in cfgs:
    m = [sum(d[(c, l)]) / max(len(d[(c, l)]), 1) for l in libs]
    print(f'{c:46s}' + ''.join(f'{x:8.1f}{x / m[0]:6.3f}' for x in m))
"
