"""Per-source-line instruction counts, lanes per instruction and stall-sample share of an .ncu-rep captured with
--import-source on (-lineinfo build). usage: ncu_source_lines.py report.ncu-rep [top] [sort column: 4 = instructions, 3 = samples]"""
import csv,sys,subprocess,io
rep=sys.argv[1]; top=int(sys.argv[2]) if len(sys.argv)>2 else 50
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--print-source','cuda,sass'],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(out)))
cur=None; res=[]
tot_i=tot_t=tot_s=0
for r in rows:
    if not r: continue
    if r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if r[0] in ('Function Name','Line No','',): continue
    try: ln=int(r[0]); samples=int(r[6]); inst=int(r[7]); tinst=int(r[8])
    except: continue
    res.append((cur,ln,r[1].strip()[:90],samples,inst,tinst)); tot_i+=inst; tot_t+=tinst; tot_s+=samples
print("total inst",tot_i,"thread inst",tot_t,"avg lanes",tot_t/tot_i,"samples",tot_s)
res.sort(key=lambda x:-x[int(sys.argv[3]) if len(sys.argv)>3 else 4])
for o in res[:top]:
    print(f"{o[0]:14s}{o[1]:4d} inst {100*o[4]/tot_i:5.1f}% lanes {o[5]/max(o[4],1):5.1f} smp {100*o[3]/tot_s:5.1f}% | {o[2]}")
