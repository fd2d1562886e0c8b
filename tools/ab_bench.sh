#!/bin/bash
# Same-box A/B of two builds of libfa_b200 (boxes differ by a few percent, so variants must share a call):
#   tools/ab_bench.sh libfa_b200_A.so libfa_b200.so [rounds]
# alternates the two libraries `rounds` times over tools/prefill_bench.py and prints the per-config means.
A=$1; B=$2; R=${3:-3}
D=vllm-flash-attention_b200/vllm_flash_attn
for r in $(seq 1 $R); do
  for v in A B; do
    lib=$A; [ $v = B ] && lib=$B
    FA_B200_LIB=$PWD/$D/$lib timeout 200 python tools/prefill_bench.py --iters 30 2>/dev/null | grep '"config"' | sed "s/^/$v /"
  done
done | python -c "
import sys, json, collections
d = collections.defaultdict(list)
for ln in sys.stdin:
    v, js = ln.split(' ', 1); o = json.loads(js); d[(o['config'], v)].append(o['TFLOPs'])
cfgs = sorted({c for c, _ in d})
for c in cfgs:
    a, b = d[(c, 'A')], d[(c, 'B')]
    ma, mb = sum(a) / len(a), sum(b) / len(b)
    print(f'{c:46s} A {ma:7.1f}  B {mb:7.1f}  B/A {mb / ma:6.3f}   (A {min(a):.0f}-{max(a):.0f}, B {min(b):.0f}-{max(b):.0f})')
"
