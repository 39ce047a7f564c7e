"""Regenerates tests/golden/eventlog_golden.{bin,npz} from the reference's own EventLog
(run in the build container only; /root/reference does not exist on the GPU box).

event_log.py cannot be imported here: it needs `event_pb2` (protoc output, git-ignored in the
reference and protoc is absent), StringIO and matplotlib.  This script
  * builds the `cp` messages of /root/reference/event.proto at run time with google.protobuf
    (a descriptor assembled from the .proto text) — the serialiser is the real protobuf one;
  * takes the SOURCE TEXT of the reference's EventLog / EventLogReader classes out of
    /root/reference/event_log.py at run time (nothing is copied into the repo) and executes it
    against that module;
  * drives the reference EventLog with a seeded scenario and stores the file it wrote plus the
    scenario's inputs.  Tests require cartpoleplusplus_b200/event_log.py to write the same
    bytes and to read them back to the same values.
"""
import os
import re
import struct
import sys
import tempfile
import types

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
TYPES = {"int32": descriptor_pb2.FieldDescriptorProto.TYPE_INT32, "float": descriptor_pb2.FieldDescriptorProto.TYPE_FLOAT,
         "bytes": descriptor_pb2.FieldDescriptorProto.TYPE_BYTES}


def event_pb2_module():
    text = re.sub(r"//[^\n]*", "", open(os.path.join(REF, "event.proto")).read())
    fd = descriptor_pb2.FileDescriptorProto(name="event.proto", package=re.search(r"package\s+(\w+);", text).group(1), syntax="proto2")
    for name, body in re.findall(r"message\s+(\w+)\s*\{(.*?)\}", text, flags=re.S):
        msg = fd.message_type.add(name=name)
        for label, typ, fname, num in re.findall(r"(optional|repeated)\s+(\w+)\s+(\w+)\s*=\s*(\d+)\s*;", body):
            f = msg.field.add(name=fname, number=int(num))
            f.label = f.LABEL_REPEATED if label == "repeated" else f.LABEL_OPTIONAL
            if typ in TYPES:
                f.type = TYPES[typ]
            else:
                f.type, f.type_name = f.TYPE_MESSAGE, "." + fd.package + "." + typ
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    mod = types.ModuleType("event_pb2")
    for m in fd.message_type:
        setattr(mod, m.name, message_factory.GetMessageClass(pool.FindMessageTypeByName(fd.package + "." + m.name)))
    return mod


def reference_classes(event_pb2):
    text = open(os.path.join(REF, "event_log.py")).read()
    start, end = text.index("class EventLog(object):"), text.index("def make_dir(d):")
    ns = {"event_pb2": event_pb2, "struct": struct, "np": np, "gzip": __import__("gzip")}
    exec(compile(text[start:end], os.path.join(REF, "event_log.py"), "exec"), ns)
    src = text[text.index("def read_state_from_event(event):"):start]
    exec(compile(src, os.path.join(REF, "event_log.py"), "exec"), ns)
    return ns["EventLog"], ns["EventLogReader"], ns["read_state_from_event"]


def scenario(seed=21):
    """episodes = list of (initial_state (R,2,7) f32, [(action, reward, state)]); action is a python int
    (discrete agents) or an array of shape (1, 2) (continuous agents)."""
    rng = np.random.default_rng(seed)
    eps = []
    for k, (R, n, discrete) in enumerate([(2, 3, False), (2, 1, False), (1, 4, True), (3, 2, False), (2, 0, False)]):
        init = rng.normal(size=(R, 2, 7)).astype(np.float32)
        seq = []
        for _ in range(n):
            a = int(rng.integers(0, 5)) if discrete else rng.uniform(-1, 1, size=(1, 2)).astype(np.float32)
            seq.append((a, 1.0, (rng.normal(size=(R, 2, 7)) * 2).astype(np.float32)))
        eps.append((init, seq))
    return eps


if __name__ == "__main__":
    pb2 = event_pb2_module()
    EventLog, EventLogReader, read_state_from_event = reference_classes(pb2)
    eps = scenario()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "log.bin")
        log = EventLog(path, False)
        for init, seq in eps:  # bullet_cartpole.py:342-344 (reset) and :272-273 (step)
            log.reset()
            log.add_just_state(init)
            for a, r, s in seq:
                log.add(s, a, r)
        log.reset()  # the reference only writes an episode when the next one starts
        log.log_file.close()
        data = open(path, "rb").read()
        # what the reference's reader makes of it
        n_events = [len(e.event) for e in EventLogReader(path).entries()]
        first = next(iter(EventLogReader(path).entries()))
        st = read_state_from_event(first.event[1])
    open(os.path.join(HERE, "eventlog_golden.bin"), "wb").write(data)
    out = {"n_episodes": np.array([len(eps)]), "events_per_episode": np.array(n_events), "first_step_state": st}
    for k, (init, seq) in enumerate(eps):
        out[f"ep{k}/init"] = init
        out[f"ep{k}/n"] = np.array([len(seq)])
        if seq:
            out[f"ep{k}/discrete"] = np.array([int(isinstance(seq[0][0], int))])
            out[f"ep{k}/actions"] = np.stack([np.asarray(a, dtype=np.float32).reshape(-1) for a, _, _ in seq])
            out[f"ep{k}/rewards"] = np.array([r for _, r, _ in seq])
            out[f"ep{k}/states"] = np.stack([s for _, _, s in seq])
    np.savez_compressed(os.path.join(HERE, "eventlog_golden.npz"), **out)
    print("written", len(data), "bytes;", n_events, "events per episode")
