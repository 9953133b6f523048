# usage: bash tools/rs_sweep.sh "<rs_probe spec>" "<R list>" "<S list>" "<C list>"
SPEC=${1:-"2048,2048,2048>3072,3072,3072:u16:0"}
for R in ${2:-2 4 8}; do for S in ${3:-2 3 4}; do for C in ${4:-3 4 6 8 12}; do
  echo -n "R=$R S=$S C=$C  "; VKTB_RS_ROWS=$R VKTB_RS_STAGES=$S VKTB_RS_CTAS=$C python tools/rs_probe.py "$SPEC" 2>&1 | awk '{print $2, $3}'
done; done; done
