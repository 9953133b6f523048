U="2048,2048,2048>3072,3072,3072:u16:0"
for i in 1 2 3; do
 echo -n "new-default "; python tools/rs_probe.py "$U" | awk '{print $2}'
 echo -n "R2S2 "; VKTB_RS_ROWS=2 VKTB_RS_STAGES=2 python tools/rs_probe.py "$U" | awk '{print $2}'
 echo -n "R3S3(old) "; VKTB_RS_ROWS=3 VKTB_RS_STAGES=3 python tools/rs_probe.py "$U" | awk '{print $2}'
 echo -n "R4S2 "; VKTB_RS_ROWS=4 VKTB_RS_STAGES=2 python tools/rs_probe.py "$U" | awk '{print $2}'
done
python tools/rs_probe.py "2048,2048,2048>3072,3072,3072:u8:0" "2048,2048,2048>1024,1024,1024:u16:0" "2048,2048,2048>1024,1024,1024:u8:0" "1024,1024,1024>1536,1536,1536:f32:0" | awk '{print $1, $2}'
