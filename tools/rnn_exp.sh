# timing ablations of the persistent recurrence (DS2_RNN_EXP bits: 1 relaxed publish, 2 no acquire wait, 4 no h loads,
# 8 no publish barrier); results are invalid when set, only the step time is of interest
for e in ${EXPS:-0 1 15}; do echo "EXP=$e"; DS2_RNN_EXP=$e timeout 120 python tools/rnn_prof.py 366 1760 32 2>&1 | grep -v "^$" | head -20; done
