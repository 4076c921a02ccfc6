"""Public-name diff against the reference: every non-test ``.py`` file of the reference is parsed (no import), its
module-level functions, classes, public methods and constants listed, and each looked up in the file of the same path
here — by AST, following ``from x import ...`` re-exports, and for methods by importing the module and asking the class
(inherited methods count).  Prints what is missing per file and a total.

    python tools/api_diff.py [reference root]
"""
import ast, importlib, os, sys
OURS = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = sys.argv[1] if len(sys.argv) > 1 else '/root/reference'
sys.path.insert(0, OURS)
def names(path):
    try: tree=ast.parse(open(path).read())
    except Exception as e: return None
    out=set()
    for n in tree.body:
        if isinstance(n,(ast.FunctionDef,ast.AsyncFunctionDef)):
            if not n.name.startswith('_'): out.add(n.name)
        elif isinstance(n,ast.ClassDef):
            if n.name.startswith('_'): continue
            out.add(n.name)
            for m in n.body:
                if isinstance(m,(ast.FunctionDef,)) and not m.name.startswith('_'): out.add(n.name+'.'+m.name)
        elif isinstance(n,ast.Assign):
            for t in n.targets:
                if isinstance(t,ast.Name) and not t.id.startswith('_'): out.add(t.id)
                elif isinstance(t,ast.Tuple):
                    for e in t.elts:
                        if isinstance(e,ast.Name) and not e.id.startswith('_'): out.add(e.id)
        elif isinstance(n,ast.AnnAssign) and isinstance(n.target,ast.Name): out.add(n.target.id)
    return out
def resolve(path):
    # follow re-export modules: collect names from 'from x import *' or import names
    tree=ast.parse(open(path).read()); out=names(path)
    for n in tree.body:
        if isinstance(n,ast.ImportFrom):
            for a in n.names:
                if a.name=='*':
                    mod=n.module.replace('.','/')
                    for cand in (os.path.join(OURS,mod+'.py'),os.path.join(OURS,mod,'__init__.py')):
                        if os.path.exists(cand): out|=resolve(cand); break
                else: out.add(a.asname or a.name)
    return out
tot=0
for root,_,files in os.walk(REF):
    if '.git' in root: continue
    for f in files:
        if not f.endswith('.py') or f.endswith('_test.py'): continue
        rp=os.path.join(root,f); rel=os.path.relpath(rp,REF)
        op=os.path.join(OURS,rel)
        rn=names(rp)
        if not rn: continue
        if not os.path.exists(op):
            print('MISSING FILE',rel,len(rn)); tot+=len(rn); continue
        on=resolve(op)
        # class methods: check by import? approximate: for 'C.m' if C re-exported, load attr
        miss=sorted(x for x in rn if x not in on)
        miss2=[]
        for x in miss:
            if '.' in x:
                c,m=x.split('.')
                if c in on:                          # inherited or re-exported: ask the class itself
                    try:
                        mod=importlib.import_module(rel[:-3].replace('/','.'))
                        if hasattr(getattr(mod,c),m): continue
                    except Exception as e: pass
            miss2.append(x)
        if miss2:
            print(rel,':',', '.join(miss2)); tot+=len(miss2)
print('total',tot)
