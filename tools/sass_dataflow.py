"""Lists the floating-point dataflow of a straight-line SASS region as non-contractable CUDA intrinsics.
Used once to derive csrc/surfel_math_pinned.cuh from the oracle build (oracle/_ref):
  cuobjdump -sass -fun <preprocessCUDA> oracle/_ref/.../_C.so | <strip encodings> > /tmp/ref_pre.sass
  python tools/sass_dataflow.py <first line> <last line>
The register->array naming (rot/scale/p/PM/VM) is specific to that kernel."""
import re,sys
lines=open('/tmp/ref_pre.sass').read().splitlines()
lo,hi=int(sys.argv[1]),int(sys.argv[2])
ver={}   # reg -> expression name
defs=[]  # (name, expr)
cnt=[0]
base={'R20':'rot','R10':'scale','R30':'p','R4':'PM'}
UR={'UR4':'W','UR5':'H'}
def new(expr):
    cnt[0]+=1; n='t%d'%cnt[0]; defs.append((n,expr)); return n
def src(tok):
    tok=tok.strip().rstrip(';').strip()
    neg=tok.startswith('-'); tok=tok.lstrip('-')
    ab=tok.startswith('|'); tok=tok.strip('|')
    tok=tok.replace('.reuse','')
    if tok=='RZ': v='0.0f'
    elif tok in ver: v=ver[tok]
    elif tok in UR: v=UR[tok]
    elif re.match(r'^-?[0-9.e+-]+$',tok): v=tok+'f'
    else: v='?'+tok
    if ab: v='fabsf(%s)'%v
    if neg: v='-'+v
    return v
for ln in lines[lo-1:hi]:
    m=re.match(r'/\*([0-9a-f]+)\*/\s+(@!?U?P\d\s+)?(\S+)\s*(.*);',ln)
    if not m: continue
    addr,pred,op,args=m.groups()
    a=[x.strip() for x in args.split(',')]
    if op.startswith('LDC.64') and a[0]=='R4':
        base['R4']='VM' if '0x3d8' in args else 'PM'; continue
    if op.startswith('LDG'):
        mm=re.search(r'\[(R\d+)\.64(\+0x([0-9a-f]+))?\]',args)
        b=mm.group(1); off=int(mm.group(3),16)//4 if mm.group(3) else 0
        ver[a[0]]='%s[%d]'%(base.get(b,b),off); continue
    if op.startswith('UIADD3'):
        UR[a[0]]='(%s-1)'%UR.get(a[3],a[3]) if a[4]=='-0x1' else '?'; continue
    if op.startswith('I2FP'):
        ver[a[0]]=new('(float)%s'%UR.get(a[1],a[1])); continue
    if op=='FMUL':
        e='__fmul_rn(%s, %s)'%(src(a[1]),src(a[2]))
    elif op=='FFMA':
        e='__fmaf_rn(%s, %s, %s)'%(src(a[1]),src(a[2]),src(a[3]))
    elif op=='FADD':
        e='__fadd_rn(%s, %s)'%(src(a[1]),src(a[2]))
    elif op=='MUFU.RSQ':
        e='rsq(%s)'%src(a[1])
    elif op in('MOV','IMAD.MOV.U32'):
        ver[a[0]]=src(a[-1]); continue
    elif op=='HFMA2':
        ver[a[0]]='1.0f' ; continue
    elif op=='FSEL':
        e='(%s ? %s : %s)'%(a[3],src(a[1]),src(a[2]))
    elif op.startswith('FSETP'):
        print('// %s %s'%(op,[src(x) if x.startswith(('R','-','|')) else x for x in a])); continue
    elif op.startswith('STG'):
        mm=re.search(r'\[(R\d+)\.64(\+0x([0-9a-f]+))?\]',args)
        off=int(mm.group(3),16)//4 if mm.group(3) else 0
        defs.append(('OUT[%d]'%off, src(a[-1]))); continue
    else:
        print('// skip',addr,op,args); continue
    if pred: e='(%s) %s else keep %s'%(pred.strip(),e,ver.get(a[0],'?'))
    ver[a[0]]=new(e)
for n,e in defs: print('%s = %s;'%(n,e))
print({k:v for k,v in ver.items() if k in('R4','R5','R6','R24')})
