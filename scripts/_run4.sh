export PF_TC_DEBUG=1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
unset PF_TC_DEBUG
for ev in 512 4096; do
for m in raw graph; do timeout 300 python scripts/prof_train.py $ev $m; done 2>&1 | grep -v "^+"
PF_TRAIN_TC=0 timeout 300 python scripts/prof_train.py $ev graph 2>&1 | grep -v "^+" | sed 's/^/[fp32 gemms] /'
PF_TRAIN_STREAMS=0 timeout 300 python scripts/prof_train.py $ev graph 2>&1 | grep -v "^+" | sed 's/^/[one stream] /'
done
