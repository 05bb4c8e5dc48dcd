export PF_TC_DEBUG=1
for v in f1 tc0 tc1; do for r in 64 128 129 300; do timeout 300 python scripts/diag_train_tc.py $r lean_default $v 2>&1 | grep -v "^  grad t" ; done; done
