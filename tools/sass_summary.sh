# SASS evidence per kernel of the built libflrl.so: sm_100a cubins, TMA bulk copies (UBLKCP / UTMALDG), mbarriers (SYNCS), wide stores.
# usage: bash tools/sass_summary.sh > profiles/r02_sass_summary.txt
LIB=fl-rl-compression-mpi_b200/libflrl.so
echo "# cuobjdump -sass $LIB | per-kernel mnemonic counts ($(date -u +%F))"
cuobjdump -lelf $LIB | sed 's/^/# /'
cuobjdump -sass $LIB | awk '
/Function :/ { fn=$3; next }
fn != "" {
  if ($0 ~ /UBLKCP/) c[fn,"UBLKCP"]++
  if ($0 ~ /UTMALDG/) c[fn,"UTMALDG"]++
  if ($0 ~ /UBLKPF/) c[fn,"UBLKPF"]++
  if ($0 ~ /SYNCS/) c[fn,"SYNCS"]++
  if ($0 ~ /STG\.E\.ENL2\.256/) c[fn,"STG.256"]++
  if ($0 ~ /STG\.E\.128/) c[fn,"STG.128"]++
  if ($0 ~ /IDP\.4A/) c[fn,"IDP.4A"]++
  if ($0 ~ /PRMT/) c[fn,"PRMT"]++
  fns[fn]=1
}
END { printf "%-90s %7s %7s %7s %6s %6s %7s %7s %6s\n","kernel","UBLKCP","UTMALDG","UBLKPF","SYNCS","STG256","STG128","IDP.4A","PRMT";
  for (f in fns) printf "%-90s %7d %7d %7d %6d %6d %7d %7d %6d\n", f, c[f,"UBLKCP"], c[f,"UTMALDG"], c[f,"UBLKPF"], c[f,"SYNCS"], c[f,"STG.256"], c[f,"STG.128"], c[f,"IDP.4A"], c[f,"PRMT"] }'
