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
        wrap.reset(seed=rank)
        consumer = torch.cuda.Stream(device=dev)
        for roots in (None, [0]):
            xchg = PeerObservationExchange((len(cams), B, H, W, 3), dev, roots=roots)
            root = roots is None or rank in roots
            for mode in ("joined", "deferred", "after"):
                for k in range(5):  # both buffers are reused
                    xchg.begin(k, wrap._engine)
                    wrap.obs_out = xchg.local_out(k)
                    wrap.obs_peers = xchg.destinations(k)
                    wrap.defer_delivery_join = {"joined": False, "deferred": True, "after": "after"}[mode]
                    wrap.sync_each_step = mode == "joined"
                    obs = wrap.step(None)[0]
                    xchg.publish(k, wrap._engine)
                    mine = torch.stack([obs["sensor_data"][c]["rgb"] for c in cams])  # (C,B,H,W,3)
                    assert obs["sensor_data"][cams[0]]["rgb"].data_ptr() == xchg.local_out(k).data_ptr()  # in place
                    want = torch.empty((world,) + tuple(mine.shape), dtype=torch.uint8, device=dev)
                    dist.all_gather_into_tensor(want.view(-1), mine.contiguous().view(-1))
                    if root:
                        xchg.wait_ready(k)
                        torch.cuda.synchronize()
                        wrap._engine.finish()
                        got = xchg.gathered(k)
                        assert torch.equal(got, want), f"rank {rank} {mode} {roots} step {k}: differs from the all-gather"
                        assert not torch.equal(want[0], want[1])  # the shards really differ
                        xchg.release(k)
                    dist.barrier()
            # a slow reader on its own stream: step k is read (after a long spin) while steps k+1, k+2 are rendered
            # and exchanged; with two buffers step k+2 rewrites the buffer being read unless release() is honoured
            snaps, wants = [], []
            for k in range(5, 11):
                xchg.begin(k, wrap._engine)
                wrap.obs_out, wrap.obs_peers = xchg.local_out(k), xchg.destinations(k)
                wrap.defer_delivery_join, wrap.sync_each_step = True, False
                obs = wrap.step(None)[0]
                if root and k > 5:  # consume step k - 1 one step behind, slowly
                    with torch.cuda.stream(consumer):
                        xchg.wait_ready(k - 1)
                        torch.cuda._sleep(int(3e6))  # ~1.5 ms of spinning before the read
                        snaps.append(xchg.gathered(k - 1).clone())
                        xchg.release(k - 1)
                xchg.publish(k, wrap._engine)
                mine = torch.stack([obs["sensor_data"][c]["rgb"] for c in cams])
                want = torch.empty((world,) + tuple(mine.shape), dtype=torch.uint8, device=dev)
                dist.all_gather_into_tensor(want.view(-1), mine.contiguous().view(-1))
                wants.append(want)
            torch.cuda.synchronize()
            for i, snap in enumerate(snaps):
                assert torch.equal(snap, wants[i]), f"rank {rank} {roots}: slow reader saw step {5 + i} overwritten"
            dist.barrier()
            del xchg
    finally:
        dist.destroy_process_group()


def test_peer_exchange_matches_all_gather():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, 29533), nprocs=2, join=True)
