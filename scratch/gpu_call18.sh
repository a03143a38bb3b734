#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c18; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu -x --durations=5 > $O/suite.log 2>&1; echo "suite rc=$?" >> $O/rc.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -25 $O/suite.log; tail -3 $O/smoke.log
