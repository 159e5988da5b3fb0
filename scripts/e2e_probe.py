"""Host-side time line of CLSM.train_on_batch under data parallelism (torchrun, cfg2): staging, issue, wait for the loss."""
import os, sys, time, argparse
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as BN
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
c = dict(BN.CFG2)
args = argparse.Namespace(conv_impl="auto", lazy_adam=False)
net = BN.build_net(c, lr, args, torch)
if world > 1:
    from dnn_page_vectors_b200.parallel import DataParallel
    DataParallel(net)
batches = [[torch.as_tensor(a).pin_memory() for a in x] for x in BN.make_batches(c, 4, 1337 + rank)]
for i in range(4):
    net.train_on_batch(batches[i % 4])
torch.cuda.synchronize()
rows = []
for i in range(8):
    t0 = time.perf_counter()
    B, bq, bd = net._stage(batches[i % 4], True)
    t1 = time.perf_counter()
    h = net.step_resident(B, bq, bd)
    t2 = time.perf_counter()
    la = h["loss_acc"].cpu()
    t3 = time.perf_counter()
    rows.append((t1 - t0, t2 - t1, t3 - t2))
r = np.median(np.array(rows), axis=0) * 1e3
print("rank %d: stage %.2f ms | issue %.2f ms | wait for loss %.2f ms | total %.2f ms" % (rank, r[0], r[1], r[2], r.sum()), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
