for r in 8 16 24 32 48; do echo "reserve $r"; HB_SESSION_RESERVE_SMS=$r timeout 600 python tools/session_trace.py train 2>&1 | grep -E "tick:"; done
