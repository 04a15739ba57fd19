import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
cur=None; out=[]; hdr=None
for r in rows:
    if len(r)==2 and r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if len(r)>6 and r[0]=='Line No': hdr=r; continue
    if hdr and len(r)>6 and r[0]!='' :
        try: s=int(r[6])
        except: continue
        stall={hdr[i][6:]:int(r[i]) for i in range(len(hdr)) if hdr[i].startswith('stall_') and 'Not' not in hdr[i] and i<len(r) and r[i].isdigit() and int(r[i])>0}
        out.append((cur,int(r[0]),r[1].strip()[:95],s,int(r[7]),sorted(stall.items(),key=lambda kv:-kv[1])[:2]))
tot=sum(o[3] for o in out)
print("total",tot)
for o in sorted(out,key=lambda o:-o[3])[:int(sys.argv[2]) if len(sys.argv)>2 else 30]:
    print(o[0][:12],o[1],o[3],o[4],o[5],o[2])
