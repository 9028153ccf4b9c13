for vec in 4 8; do for ch in 2 4; do for per in 24 48; do for th in 96 192 384; do
echo "VEC=$vec CH=$ch PER=$per TH=$th"; WVB200_AGG_VEC=$vec WVB200_AGG_CHUNK=$ch WVB200_AGG_PER=$per WVB200_AGG_THREADS=$th python tools/aggtime.py kernel 2>&1
done; done; done; done
