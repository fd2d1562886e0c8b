#!/bin/bash
# Which Blackwell instructions the built library contains (no GPU needed): tcgen05 MMAs (UTCHMMA), TMEM loads / stores
# (LDTM / STTM), TMA tensor loads / stores (UTMALDG / UTMASTG), MMA commit barriers (UTCBAR), mbarrier ops (SYNCS),
# register re-allocation (USETMAXREG), packed fp32 math (FFMA2 / FADD2 / FMUL2) — and, for the decode / tile kernels,
# mma.sync (HMMA.16816) with cp.async (LDGSTS).  Per kernel family.
#   bash tools/sass_evidence.sh > profiles/r02_sass_counts.txt
LIB=${1:-vllm-flash-attention_b200/vllm_flash_attn/libfa_b200.so}
PAT='\b(UTCHMMA|LDTM|STTM|UTMALDG|UTMASTG|UTCBAR|HMMA\.16816|LDGSTS|MUFU\.EX2|SYNCS|USETMAXREG|FFMA2|FADD2|FMUL2|STL|LDL)[A-Za-z0-9_.]*'
echo "# $(basename $LIB): SASS mnemonic counts per kernel family (cuobjdump -sass), nvcc $(nvcc --version | grep -o 'release [0-9.]*')"
for fam in fa_prefill_tc5_kernel fa_decode_kernel fa_tile_kernel fa_combine_kernel fa_exchange_combine_kernel fa_append_kernel tc5_schedule_kernel; do
  echo "== $fam"
  cuobjdump -sass "$LIB" 2>/dev/null | awk -v fam="$fam" '/Function :/{on = index($0, fam) > 0} on' | grep -oE "$PAT" | sed -E 's/\.(E|BYPASS|128|64|F32|BF16|x16|x32|4D|IM2COL|CTAPOOL|TRANS64|A1T0|EXCH|PHASECHK|TRYWAIT|ARRIVE|DEALLOC|TRY_ALLOC|2D|3D|5D|RECONVERGENT)\b//g' | sort | uniq -c | sort -rn | head -14
done
