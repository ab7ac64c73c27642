for v in 3 4; do
for c in 740 2960; do
for it in sb,pc,ps,ts,vdw,hb,hp sb,pc,ps,ts,vdw,hb; do
echo "variant $v chunk $c itypes $it"
CHACRA_BIN_VARIANT=$v CHACRA_CHUNK_FRAMES=$c python tools/profile_run.py --frames 2960 --launches 3 --itypes $it 2>&1 | tail -1
done; done; done
