for mode in 0 2; do for g in 1 2 4; do
echo "== OUT_MODE=$mode GROUP=$g"; SMSGATE_OUT_MODE=$mode SMSGATE_D2H_GROUP=$g python scratch/timeline.py 2>&1 | tail -n 4
done; done
