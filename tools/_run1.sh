python tools/zero_check1.py 2>&1 | grep -v Warn | tail -6
