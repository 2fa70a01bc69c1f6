python tools/tc_dbg.py 2>&1 | tail -4
