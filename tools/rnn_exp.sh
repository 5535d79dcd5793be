# Timing ablations / variants of the persistent recurrence, selected with DS2_RNN_EXP (a bit mask; see the p.exp uses in
# deepspeech_b200/csrc/ds2_rnn_persistent.cu).  Bits 2, 4, 8 give INVALID results - only the step time is of interest:
#   2  no wait on the step counter        4  no loads of h_{t-1}          8  no CTA barrier before the publication
#  16  odd instances start 1.8 us late   32  direction 1 starts 0.9 us late          (stagger experiments, valid results)
# 128  fp32 partial sums through one 2 KB cp.async.bulk per warp (valid)   256  ld.global.cg instead of ld.relaxed.gpu (valid)
# Related environment switches: DS2_RNN_FENCED=1 (release/acquire counters), DS2_RNN_SLAB_F32=1 (fp32 slab exchange).
for e in ${EXPS:-0 14}; do echo "EXP=$e"; DS2_RNN_EXP=$e timeout 120 python tools/rnn_prof.py 366 1760 32 2>&1 | grep -v "^$" | head -20; done
