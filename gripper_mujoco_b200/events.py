"""Host mirrors of MjType::EventTrack / MjType::TestReport (src/mjclass.h:575-660, 941-950; bound in
src/bind.cpp:367-560): one attribute per event name, each with value / last_value / active_sum / row / abs / percent.

The counters themselves live on the GPU (EnvState::ev_row / ev_abs / ev_last); gmj_event_track_host copies them out
and `event_tracks()` below wraps them, one EventTrack per environment.
"""
from __future__ import annotations

from .settings import BINARY_NAMES, LINEAR_NAMES


class BinaryEvent:
    """MjType::EventTrack::BinaryEvent (src/mjclass.h:577-586)."""
    __slots__ = ("value", "last_value", "active_sum", "row", "abs", "percent")

    def __init__(self):
        self.reset()

    def reset(self):
        self.value = False
        self.last_value = False
        self.active_sum = 0
        self.row = 0
        self.abs = 0
        self.percent = 0.0

    def __repr__(self):
        return f"BinaryEvent(row={self.row}, abs={self.abs}, last_value={self.last_value}, percent={self.percent:.1f})"


class LinearEvent(BinaryEvent):
    """MjType::EventTrack::LinearEvent (src/mjclass.h:588-597): value / last_value are floats."""
    __slots__ = ()

    def reset(self):
        super().reset()
        self.value = 0.0
        self.last_value = 0.0


class EventTrack:
    """MjType::EventTrack: `track.lifted.abs`, `track.palm_force.last_value`, ... (X-macro order of simsettings.h)."""

    def __init__(self):
        for nm in BINARY_NAMES:
            setattr(self, nm, BinaryEvent())
        for nm in LINEAR_NAMES:
            setattr(self, nm, LinearEvent())

    def reset(self):                                   # src/mjclass.h:613-627
        for nm in BINARY_NAMES + LINEAR_NAMES:
            getattr(self, nm).reset()

    def calculate_percentage(self):                    # src/mjclass.h:629-641: percent of steps with the event
        steps = float(self.step_num.abs)
        for nm in BINARY_NAMES + LINEAR_NAMES:
            ev = getattr(self, nm)
            ev.percent = (100.0 * ev.abs) / steps if steps > 0 else 0.0

    def print(self):                                   # src/mjclass.h:643-660
        print(str(self))

    def __str__(self):
        rows = ["EventTrack: name (row, abs, last_value)"]
        for nm in BINARY_NAMES + LINEAR_NAMES:
            ev = getattr(self, nm)
            rows.append(f"  {nm:28s} ({ev.row}, {ev.abs}, {ev.last_value})")
        return "\n".join(rows)

    def __getstate__(self):                            # src/bind.cpp:547-560 pickles every event
        return {nm: {k: getattr(getattr(self, nm), k) for k in BinaryEvent.__slots__} for nm in BINARY_NAMES + LINEAR_NAMES}

    def __setstate__(self, st):
        self.__init__()
        for nm, d in st.items():
            ev = getattr(self, nm)
            for k, v in d.items():
                setattr(ev, k, v)


class TestReport:
    """MjType::TestReport (src/mjclass.h:941-950): what MjEnv._monitor_test reads after an episode."""
    __test__ = False               # not a pytest class

    def __init__(self, object_name="", cumulative_reward=0.0, cnt=None):
        self.object_name = object_name
        self.cumulative_reward = cumulative_reward
        self.cnt = cnt if cnt is not None else EventTrack()


def add_events(e1: EventTrack, e2: EventTrack) -> EventTrack:
    """MjClass::add_events (src/mjclass.cpp:4692-4715): abs and active_sum add; linear events also add last_value;
    everything else is left at its reset value."""
    out = EventTrack()
    for nm in BINARY_NAMES:
        a, b, o = getattr(e1, nm), getattr(e2, nm), getattr(out, nm)
        o.abs = a.abs + b.abs
        o.active_sum = a.active_sum + b.active_sum
    for nm in LINEAR_NAMES:
        a, b, o = getattr(e1, nm), getattr(e2, nm), getattr(out, nm)
        o.abs = a.abs + b.abs
        o.active_sum = a.active_sum + b.active_sum
        o.last_value = a.last_value + b.last_value
    return out


def event_tracks(row, abs_, last, env_ids=None):
    """EventTrack objects from gmj_event_track_host's arrays ([n, 36] each, binary events first)."""
    names = BINARY_NAMES + LINEAR_NAMES
    nb = len(BINARY_NAMES)
    out = []
    for e in (range(row.shape[0]) if env_ids is None else env_ids):
        t = EventTrack()
        for k, nm in enumerate(names):
            ev = getattr(t, nm)
            ev.row = int(row[e, k]); ev.abs = int(abs_[e, k])
            ev.active_sum = int(row[e, k] > 0)                         # src/mjclass.cpp:5448 / 5460
            ev.last_value = bool(last[e, k] != 0) if k < nb else float(last[e, k])
            # `value` was consumed into last_value and cleared by update_events (src/mjclass.cpp:5442-5460)
        out.append(t)
    return out
