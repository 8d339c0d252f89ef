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
