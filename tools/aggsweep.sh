for ch in 2 4; do for per in 8 24 100; do for smax in 2 4 8; do for th in 192 384 512; do
echo "CH=$ch PER=$per SMAX=$smax TH=$th"; WVB200_AGG_CHUNK=$ch WVB200_AGG_PER=$per WVB200_AGG_SMAX=$smax WVB200_AGG_THREADS=$th python tools/aggtime.py kernel 2>&1
done; done; done; done
