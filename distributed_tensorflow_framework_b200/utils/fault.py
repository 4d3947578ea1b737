"""Fault injection for recovery tests (the reference has none): DTF_KILL_RANK_AT_STEP="<rank>:<step>" makes
that worker die abruptly when it reaches the step, like a preempted task. Surviving ranks detect it through the
heartbeat hook (runtime/hooks.py) or the collective's bounded spin, and the job is restarted from the last checkpoint."""
import os


def parse(spec=None):
    spec = os.environ.get("DTF_KILL_RANK_AT_STEP", "") if spec is None else spec
    if not spec:
        return None
    try:
        r, s = spec.split(":")
        return int(r), int(s)
    except ValueError:
        raise ValueError("DTF_KILL_RANK_AT_STEP must look like '<rank>:<step>', got %r" % spec)


def maybe_kill(rank, step, spec=None, _exit=os._exit):
    target = parse(spec)
    if target is not None and target == (int(rank), int(step)):
        _exit(17)
        return True
    return False
