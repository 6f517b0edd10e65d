"""Multi-GPU check of the peer exchange, run under torchrun on a box with >= 2 GPUs (not collected by pytest):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/peer_exchange_check.py
Each rank renders its own views of a small scene; the SH gradient formed (a) from factors read out of peer memory
(PeerFactorExchange), (b) from NCCL-all-gathered factors (finish_sh_gradients) and (c) by all-reducing the dense dL_dsh
must agree on every rank, for several steps (the two slots alternate); the other gradients summed by the two-shot
peer-memory all-reduce must match NCCL's and be bit-identical on all ranks.  Prints PEER_EXCHANGE_OK on rank 0."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from helpers import R, rel_err, settings_for, syn  # noqa: E402
from langsplat_b200 import dist as D  # noqa: E402

NAMES = ("means3D", "shs", "opacities", "scales", "rotations", "language_feature")


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    P, per_rank = 30000, 2
    scene, cams = syn.make_config(4, P=P, W=256, H=192, n_cameras=world * per_rank)
    mine = [rank * per_rank + i for i in range(per_rank)]  # rank-major = the order the exchanges use

    def run(sink):
        params = {k: scene[k].to(dev).requires_grad_(True) for k in NAMES}
        means2D = torch.zeros_like(params["means3D"], requires_grad=True)

        def views():
            for v in mine:
                cam = cams[v]
                dcol, dfeat = [t.to(dev) for t in syn.make_grads(cam.image_width, cam.image_height, 300 + v)]
                rast = R.GaussianRasterizer(settings_for(cam, dev, sh_degree=3))
                color, feat, radii = rast(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                          language_feature_precomp=params["language_feature"],
                                          opacities=params["opacities"], scales=params["scales"],
                                          rotations=params["rotations"])
                torch.autograd.backward([color, feat], [dcol, dfeat])
        others = [k for k in NAMES if k != "shs"]
        if sink == "dense":
            views()
            D.allreduce_gradients([params[k].grad for k in NAMES])
            return params["shs"].grad, [params[k].grad for k in others]
        with D.factored_sh_gradients(sink.begin_step() if sink else None) as factors:
            views()
        if sink:
            sh, red = sink.exchange(params["means3D"], 16, [params[k].grad for k in others] + [None])
            assert red[-1] is None
            return sh, [t.clone() for t in red[:-1]]
        D.allreduce_gradients([params[k].grad for k in others])
        return D.finish_sh_gradients(factors, params["means3D"], 16), [params[k].grad for k in others]

    dense, dense_others = run("dense")
    gathered, _ = run(None)
    assert rel_err(gathered, dense) <= 1e-5, rel_err(gathered, dense)
    ex = D.make_peer_exchange(P, per_rank, dev, grad_floats_per_gaussian=14)
    assert ex is not None, "peer memory could not be mapped between the ranks"
    for step in range(4):
        got, others = run(ex)
        assert rel_err(got, dense) <= 1e-5, (step, rel_err(got, dense))
        assert torch.equal(got, gathered) or rel_err(got, gathered) <= 1e-6
        for a, b in zip(others, dense_others):  # two-shot all-reduce over peer memory vs NCCL (+ atomics noise)
            assert a.shape == b.shape and rel_err(a, b) <= 2e-5, (step, rel_err(a, b))
        # every rank must hold the same bits: each chunk is reduced by one rank and broadcast
        flat = torch.cat([t.reshape(-1) for t in others])
        ref = flat.clone()
        dist.broadcast(ref, src=0)
        assert torch.equal(flat, ref), "ranks disagree on the reduced gradients"
    ex.check()
    ex.close()
    dist.barrier()
    if rank == 0:
        print("PEER_EXCHANGE_OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
