"""Diagnostic: does b200env_debug_stream_gate hold a stream in a fresh process, with and without a device
synchronisation beforehand (the pytest run of round 2 saw it fall through once)?"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from madrona_rl_envs_playground_b200 import stream_gate

for presync in (False, True):
    stream = torch.cuda.Stream()
    flag = torch.zeros(1, dtype=torch.int32).pin_memory()
    x = torch.zeros(1, device="cuda")
    if presync:
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    stream_gate(0, stream, flag, 3000)
    with torch.cuda.stream(stream):
        x += 1
    e1.record(stream)
    time.sleep(0.05)
    held = not e1.query()
    flag.fill_(1)
    stream.synchronize()
    print(f"presync={presync}: held after 50 ms = {held}, event window {e0.elapsed_time(e1):.2f} ms, x = {x.item()}", flush=True)
