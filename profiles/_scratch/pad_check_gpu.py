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
with graphed_tail.no_capture():
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
# finer: sub-module outputs of layer 0
from hydragnn.ops import node as nd
layer=model.graph_convs[0]
d_a=mk(); d_b=static["data"]
with graphed_tail.no_capture(), torch.no_grad():
    xa,_,ca=model._embedding(d_a); ca["cu_seqlens"]=model._attention_segments(d_a, xa.size(0), xa.device)
    xb,_,cb=model._embedding(d_b); cb["cu_seqlens"]=model._attention_segments(d_b, xb.size(0), xb.device)
    print("embed diff", float((xa-xb[:N]).abs().max()))
    ha,_=layer.conv(inv_node_feat=xa, equiv_node_feat=d_a.pos, **{k:v for k,v in ca.items() if k!="cu_seqlens"})
    hb,_=layer.conv(inv_node_feat=xb, equiv_node_feat=d_b.pos, **{k:v for k,v in cb.items() if k!="cu_seqlens"})
    print("local conv diff", float((ha-hb[:N]).abs().max()), float(ha.abs().max()))
    t1a,qa=layer._tail_pre(xa,ha); 
    nd.valid_rows=static["valid_rows"]; t1b,qb=layer._tail_pre(xb,hb)
    print("t1 diff", float((t1a-t1b[:N]).abs().max()), "qkv diff", float((qa-qb[:N]).abs().max()))
    from hydragnn.ops import attn_varlen
    C=layer.channels
    oa=attn_varlen(qa[:,:C],qa[:,C:2*C],qa[:,2*C:],ca["cu_seqlens"],layer.heads)
    ob=attn_varlen(qb[:,:C],qb[:,C:2*C],qb[:,2*C:],cb["cu_seqlens"],layer.heads)
    print("attn diff", float((oa-ob[:N]).abs().max()), "cu", cb["cu_seqlens"].tolist())
    nd.valid_rows=None; ya=layer._tail_post(xa,t1a,oa)
    nd.valid_rows=static["valid_rows"]; yb=layer._tail_post(xb,t1b,ob); nd.valid_rows=None
    print("post diff", float((ya-yb[:N]).abs().max()))
