"""Multi-GPU observation exchange through peer memory (gsw_frames_desc.peer_rgb_u8 + dist.PeerObservationExchange)
against the NCCL all-gather of the same frames.  Needs two GPUs on the box (skipped otherwise): run with
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_peer_exchange.py`."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port):
    import torch.distributed as dist
    from gsworld_b200 import fixtures as fx
    from gsworld_b200.dist import PeerObservationExchange
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        assert PeerObservationExchange.available()
        B, H, W = 5, 64, 48
        objects = ("006_mustard_bottle",)
        scene = fx.make_scene(fx.SceneSpec(robot="fr3_umi", n_total=20_000, per_link=200, objects=objects,
                                           per_object=900, seed=5)).to(dev)
        env = fx.FakeEnv("fr3_umi", num_envs=B, hw=(H, W), objects=objects, seed=100 + rank, device=dev)  # own shard
        scan = fx.forward_kinematics("fr3_umi", fx.rt.ROBOT_SCAN_QPOS["fr3_umi"][None])[0]
        wrap = GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, merged_model=scene,
                              scan_link_mats=list(scan), strict_reference=False)
        cams = list(env.cameras)
        xchg = PeerObservationExchange((len(cams), B, H, W, 3), dev)
        wrap.reset(seed=rank)
        for mode in ("joined", "deferred"):
            for k in range(5):  # both buffers are reused
                xchg.begin(k)
                wrap.obs_out = xchg.local_out(k)
                wrap.obs_peers = xchg.destinations(k)
                wrap.defer_delivery_join = mode == "deferred"
                wrap.sync_each_step = mode != "deferred"
                obs = wrap.step(None)[0]
                xchg.publish(k, wrap._engine)
                xchg.wait_ready(k)
                mine = torch.stack([obs["sensor_data"][c]["rgb"] for c in cams])  # (C,B,H,W,3)
                assert obs["sensor_data"][cams[0]]["rgb"].data_ptr() == xchg.local_out(k).data_ptr()  # in place
                want = torch.empty((world,) + tuple(mine.shape), dtype=torch.uint8, device=dev)
                dist.all_gather_into_tensor(want.view(-1), mine.contiguous().view(-1))
                torch.cuda.synchronize()
                wrap._engine.finish()
                got = xchg.gathered(k)
                assert torch.equal(got, want), f"rank {rank} {mode} step {k}: peer exchange differs from the all-gather"
                assert not torch.equal(want[0], want[1])  # the shards really differ
                dist.barrier()  # nobody overwrites buffer k before everyone has compared it
    finally:
        dist.destroy_process_group()


def test_peer_exchange_matches_all_gather():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, 29533), nprocs=2, join=True)
