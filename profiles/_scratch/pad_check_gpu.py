import sys, os, copy, numpy as np, torch
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0]=[ROOT, os.path.join(ROOT,'hydragnn-gps-support_b200'), os.path.join(ROOT,'tests')]
from hydragnn.graph import Data
from hydragnn.models import create_model
from hydragnn.utils import synthetic as syn
from hydragnn.train.step_graph import BucketedStepGraph
from hydragnn.ops import node as node_ops
from hydragnn.train.train_validate_test import get_head_indices
from hydragnn.globalAtt import graphed_tail
torch.backends.cuda.matmul.allow_tf32=False
workload=sys.argv[1]
spec=syn.WORKLOADS[workload]
rng=np.random.default_rng(3)
graphs=[spec["sample"](rng) for _ in range(8)]
batch=syn.finish_edges(syn.collate(graphs))
arch=dict(spec["arch"], num_conv_layers=int(sys.argv[2]) if len(sys.argv)>2 else 2)
torch.manual_seed(0)
model=create_model(**arch).cuda()
for m in model.modules():
    if isinstance(m, torch.nn.Dropout): m.p=0.0
    elif hasattr(m,"attn") and isinstance(getattr(m,"dropout",None),float): m.dropout=0.0
model.train()
dev={k:(v.cuda() if torch.is_tensor(v) else v) for k,v in batch.items() if k not in ("cell","pbc")}
mk=lambda: Data(**{k:(v.clone() if torch.is_tensor(v) else v) for k,v in dev.items()})
with graphed_tail.eager():
    d=mk(); pred=model(d); loss=model.loss(pred, d.y, get_head_indices(model,d))[0]
    print("eager loss", float(loss.detach()))
    class Opt: param_groups=[{"capturable":True}]; state={}
    sg=BucketedStepGraph(model, Opt(), node_bucket=32, edge_bucket=256)
    d2=mk(); key=sg._key_of(d2); padded=sg._pad(d2,key); static=sg._make_static(d2,padded,key)
    print("key", key, "N", d2.pos.size(0), "E", d2.edge_index.size(1))
    # hook intermediate activations: per conv layer output on the real rows
    acts={}
    def hook(name):
        def f(mod, inp, out):
            acts.setdefault(name, []).append(out[0].detach() if isinstance(out,(tuple,list)) else out.detach())
        return f
    hs=[c.register_forward_hook(hook(f"conv{i}")) for i,c in enumerate(model.graph_convs)]
    pred=model(mk())
    node_ops.valid_rows=static["valid_rows"]
    pred2=model(static["data"])
    node_ops.valid_rows=None
    loss2,_=sg._masked_loss(pred2, static)
    print("padded loss (no graph)", float(loss2.detach()))
    N=d2.pos.size(0)
    for k,v in acts.items():
        a,b=v[0],v[1][:N]
        print(k, "max abs diff on real rows", float((a-b).abs().max()), "scale", float(a.abs().max()))
    print("pred diff", float((pred[0]-pred2[0][:pred[0].size(0)]).abs().max()))
