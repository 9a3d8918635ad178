"""Instruction mix of the loops that contain MUFU in the named kernels of an object file: python tools/sass_loops.py <file.o|.so> <substring of the mangled kernel name>...
(static evidence for profiles/: packed FP32x2, MUFU, MOV, LDS counts per loop body)."""
import re,sys,subprocess
obj=sys.argv[1]
names=subprocess.run(['cuobjdump','-sass',obj],capture_output=True,text=True).stdout
funcs=re.findall(r'Function : (\S+)',names)
for pat in sys.argv[2:]:
    fn=[f for f in funcs if pat in f][0]
    txt=subprocess.run(['cuobjdump','-sass','-fun',fn,obj],capture_output=True,text=True).stdout
    ins=[]
    for m in re.finditer(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);',txt):
        ins.append((int(m.group(1),16),m.group(2)))
    addrs=[a for a,_ in ins]
    loops=[]
    for a,t in ins:
        m=re.search(r'BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)',t)
        if m:
            tgt=int(m.group(1),16)
            if tgt<a: loops.append((tgt,a))
    print('==',pat)
    for lo,hi in loops:
        body=[t for a,t in ins if lo<=a<=hi]
        nm=sum('MUFU' in t for t in body)
        if nm==0: continue
        def c(p): return sum(bool(re.search(p,t)) for t in body)
        print(f'  loop {lo:#x}-{hi:#x}: n={len(body)} packed={c("FFMA2|FMUL2|FADD2")} mufu={nm} mov={c(r"(^|\s)MOV |IMAD.MOV")} fmnmx={c("FMNMX")} lds={c("LDS")} fsetp={c("FSETP")} ffma={c(r"FFMA ")}')
