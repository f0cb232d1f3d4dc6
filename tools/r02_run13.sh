timeout 600 python tools/session_trace.py eval 2>&1 | grep -v Warn | tail -6
