import os, sys, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from cartpoleplusplus_b200.replay import DeviceReplayMemory, _p
from cartpoleplusplus_b200 import _lib
B, L = 65536, 200
rm = DeviceReplayMemory(1 << 24, (2, 2, 7), 2, num_envs=B, max_episode_len=L, device=0)
obs = torch.randn(B, 2, 2, 7, device="cuda"); act = torch.rand(B, 2, device="cuda"); rew = torch.ones(B, device="cuda")
done = torch.zeros(B, dtype=torch.uint8, device="cuda"); reason = torch.zeros(B, dtype=torch.uint8, device="cuda")
rm.begin(obs)
lib = _lib.load()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
def raw():
    lib.bcp_replay_observe(rm._h, _p(act), 0, _p(obs), _p(rew), _p(done), _p(reason), st)
for name, fn in (("raw C-ABI call", raw), ("python observe()", lambda: rm.observe(act, obs, rew, done, reason))):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(100): fn()
    e1.record(); torch.cuda.synchronize()
    print(name, "us per call (100 back to back, no finished episodes):", e0.elapsed_time(e1) * 10)
    rm.begin(obs)
# with finishing episodes: every env finishes every 4th call
done1 = torch.ones(B, dtype=torch.uint8, device="cuda")
rm.begin(obs); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for t in range(100):
    lib.bcp_replay_observe(rm._h, _p(act), 0, _p(obs), _p(rew), _p(done1 if t % 4 == 3 else done), _p(reason), st)
    if t % 4 == 3: lib.bcp_replay_begin(rm._h, _p(obs), None, st)
e1.record(); torch.cuda.synchronize()
print("us per call, 65536 four-step episodes finishing every 4th call:", e0.elapsed_time(e1) * 10, rm.info())
