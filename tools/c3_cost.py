#!/usr/bin/env python
"""Where a config-3 motion-primitive step spends its host time (microseconds, median).  usage: python tools/c3_cost.py"""
import io, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from morphablegraphs_b200 import synthetic as syn, MotionPrimitiveGenerator, graph_synthesis as GS
from morphablegraphs_b200.motion_state_graph import MotionStateGraphLoader, graph_to_bytes
from morphablegraphs_b200.skeleton import align_frames
ld = MotionStateGraphLoader(); ld.set_data_source(io.BytesIO(graph_to_bytes(syn.make_graph_data()))); graph = ld.build(resident=True)
gen = MotionPrimitiveGenerator(None, {"n_random_samples": 1000000, "use_local_coordinates": False}, skeleton=graph.skeleton, seed=21)
pi = GS.read_input(syn.EXAMPLE_INPUT); steps = GS.plan_steps(graph, pi); prev = GS.initial_frames(graph, pi, steps[0]["node_key"])
node = graph.nodes[steps[1]["node_key"]]
def med(fn, reps=100):
    out = []
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize(); out.append(1e6 * (time.perf_counter() - t0))
    return sorted(out)[len(out) // 2], r
t, specs = med(lambda: GS.step_specs(graph, steps[1], prev)); print("step_specs (host FK)            %8.1f us" % t)
dm = node.motion_primitive.device_model(graph.skeleton)
t, pp = med(lambda: dm.fk(np.ascontiguousarray(prev[-2:], dtype=np.float32), GS.POSE_JOINTS).cpu().numpy()); print("K4 fk of 2 frames + D2H         %8.1f us" % t)
t, specs = med(lambda: GS.step_specs(graph, steps[1], prev, pp)); print("step_specs (given positions)    %8.1f us" % t)
t, mpc = med(lambda: GS.build_constraints(graph, specs)); print("build_constraints               %8.1f us" % t)
t, s = med(lambda: gen.generate_constrained_sample(node, GS.build_constraints(graph, specs), "", prev, None)); print("build + generate_constrained_sample %4.1f us" % t)
t, sp = med(lambda: node.back_project(s, use_time_parameters=False)); print("back_project                    %8.1f us" % t)
t, fr = med(lambda: sp.get_motion_vector()); print("get_motion_vector               %8.1f us" % t)
t, fa = med(lambda: align_frames(graph.skeleton, fr, prev)); print("align_frames                    %8.1f us" % t)
synth = GS.GraphWalkSynthesizer(graph, gen)
t, _ = med(lambda: synth.run(syn.EXAMPLE_INPUT), 5); print("whole synthesis (17 steps)      %8.1f us = %.1f us per step" % (t, t / 17))
