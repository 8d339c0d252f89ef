"""Test double with the Engine interface (submit/step/drain/cancel/close_session/stats): emits a
canned reply one token per step.  Lives in tests/ only — the product has no such fallback."""
import threading
from types import SimpleNamespace


class FakeEngine:
    def __init__(self, tokenizer, reply="Hello from the fake engine.", max_context=4096, max_batch=64):
        self.tok = tokenizer
        self.reply = reply
        self.limits = SimpleNamespace(max_context=max_context, max_batch=max_batch)
        self.lock = threading.Lock()
        self.reqs = {}
        self.out = []
        self.next = 0
        self.submitted = []          # (session, tokens, kwargs)
        self.closed = []
        self.steps = 0

    def submit(self, session, tokens, temperature=0.0, top_k=0, top_p=1.0, max_new_tokens=256, seed=0, ignore_eos=False):
        with self.lock:
            rid = self.next
            self.next += 1
            reply = self.tok.encode(self.reply(tokens) if callable(self.reply) else self.reply)
            self.reqs[rid] = {"todo": list(reply), "max_new": max_new_tokens, "n": 0}
            self.submitted.append((session, list(tokens), dict(temperature=temperature, top_k=top_k, top_p=top_p,
                                                              max_new_tokens=max_new_tokens)))
            return rid

    def step(self):
        with self.lock:
            self.steps += 1
            n = 0
            for rid, r in list(self.reqs.items()):
                if r["todo"]:
                    t = r["todo"].pop(0)
                    r["n"] += 1
                    n += 1
                    if r["n"] >= r["max_new"]:
                        self.out.append((rid, t, "length")); del self.reqs[rid]
                    else:
                        self.out.append((rid, t, None))
                else:
                    self.out.append((rid, self.tok.eot, "stop")); del self.reqs[rid]
                    n += 1
            return n

    def drain(self):
        with self.lock:
            o, self.out = self.out, []
            return o

    def cancel(self, rid):
        with self.lock:
            if rid in self.reqs:
                del self.reqs[rid]
                self.out.append((rid, -1, "cancelled"))
                return True
            return False

    def close_session(self, session):
        self.closed.append(session)
        return True

    def stats(self):
        return {"steps": self.steps, "active_requests": len(self.reqs)}


class FakeLoopEngine(FakeEngine):
    """The same double with a stepping thread of its own, like ft_engine_start / ft_engine_wait / ft_engine_stop: the
    runner must never call step() while it runs."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.cv = threading.Condition(self.lock)
        self.loop_thread, self.loop_on, self.loop_stop = None, False, False
        self.stepped_by_owner = 0
        self.starts = 0

    def step(self):
        if self.loop_on and threading.current_thread() is not self.loop_thread:
            self.stepped_by_owner += 1
            raise RuntimeError("ft_engine_step: the engine runs its own stepping thread")
        n = super().step()
        with self.cv:
            self.cv.notify_all()
        return n

    def _loop(self):
        import time
        while not self.loop_stop:
            if self.reqs:
                self.step()
                time.sleep(0.0005)
            else:
                time.sleep(0.001)

    def start_loop(self):
        if self.loop_on:
            return
        self.starts += 1
        self.loop_stop, self.loop_on = False, True
        self.loop_thread = threading.Thread(target=self._loop, daemon=True)
        self.loop_thread.start()

    def stop_loop(self):
        self.loop_stop = True
        if self.loop_thread is not None:
            self.loop_thread.join(timeout=5)
        self.loop_on = False

    def wait_events(self, timeout_ms=50):
        with self.cv:
            if not self.out and self.loop_on:
                self.cv.wait(timeout_ms / 1000.0)
            return len(self.out)

    def cancel(self, rid):
        r = super().cancel(rid)
        with self.cv:
            self.cv.notify_all()
        return r
