O=gpurun_out; T=r2r
mkdir -p $O
for v in 0 1 0 1; do echo "CTG_NTT_LEGACY=$v"; CTG_NTT_LEGACY=$v REPS=8 timeout 300 python tools/run_once.py prove 22 0 2>&1 | grep total_ms | awk -F"'total_ms': " '{print $2}' | tr -d '}' | sort -n | tr '\n' ' '; echo; done
for v in 0 1; do echo "CTG_NTT_LEGACY=$v 2^20"; CTG_NTT_LEGACY=$v REPS=8 timeout 300 python tools/run_once.py prove 20 0 2>&1 | grep total_ms | awk -F"'total_ms': " '{print $2}' | tr -d '}' | sort -n | tr '\n' ' '; echo; done
