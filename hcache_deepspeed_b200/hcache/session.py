"""Multi-turn session manager: the glue between the ragged engine and the hidden-state store.

One turn of a session = (restore its context if it is not resident) -> prefill the new tokens -> decode ->
save the state the planner wants (hidden states for some layers, KV pages for others, nothing for a recomputed
prefix) -> release the KV blocks. The reference leaves all of this to its caller (DeepSpeed-MII); it is here so
the framework serves multi-turn traffic stand-alone and so benchmarks exercise the same code a deployment would.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional

import torch

from ..inference.sampling import sample_next
from .planner import LinkModel, RestorePlan, measure_link, plan_restore
from .store import HiddenStateStore, SessionState


class HCacheSessionManager:

    def __init__(self, engine, store: Optional[HiddenStateStore] = None, link: Optional[LinkModel] = None,
                 allow_recompute: bool = True, allow_kv: bool = True, measure_host_link: bool = False) -> None:
        """``measure_host_link``: calibrate the planner with the host link as it is when every rank transfers at once
        (collective over the default process group when one is initialised) instead of the built-in single-GPU rate."""
        self.engine = engine
        self.store = store or HiddenStateStore()
        if link is None and measure_host_link:
            link = measure_link(engine.model().device, tp_size=engine.model().tp_size)
        self.link = link or LinkModel(tp_size=engine.model().tp_size)
        self.allow_recompute, self.allow_kv = allow_recompute, allow_kv
        self._live_latents: Dict[int, List[torch.Tensor]] = {}
        self._live_tokens: Dict[int, List[torch.Tensor]] = {}
        self._resident = set()

    # ---- planning ---------------------------------------------------------------------------------------------
    def plan_for(self, n_tokens: int, from_nvme: bool = False) -> RestorePlan:
        spec = self.engine.model().spec
        return plan_restore(spec, n_tokens, self.link, from_nvme=from_nvme, allow_recompute=self.allow_recompute,
                            allow_kv=self.allow_kv)

    # ---- a turn -------------------------------------------------------------------------------------------------
    def is_resident(self, uid: int) -> bool:
        return uid in self._resident

    def restore(self, uid: int) -> int:
        """Bring a saved session back into the KV cache. Returns the number of restored tokens."""
        if uid in self._resident:
            return 0
        st = self.store.load(uid)
        plan = self.plan_for(st.n_tokens)
        self.engine.restore_session(uid, st.tokens, hidden=st.hidden, hidden_layers=st.layers_of("hidden"), kv=st.kv,
                                    kv_layers=st.layers_of("kv"), kv_layer_major=st.kv_layer_major,
                                    recompute_layers=st.layers_of("none"),
                                    interleave=plan.interleave or 2.5)
        self._resident.add(uid)
        # the hidden states of restored layers are already known: carry them into the next save
        self._live_tokens[uid] = [st.tokens]
        self._live_latents[uid] = [("state", st)]
        return st.n_tokens

    def prefill(self, uid: int, tokens: torch.Tensor) -> torch.Tensor:
        """Feed new tokens of ``uid`` (restoring it first if needed); returns last-token logits ``[1, vocab]``."""
        if uid not in self._resident and uid in self.store:
            self.restore(uid)
        logits, lat = self.engine.put([uid], [tokens], capture_latents=True)
        self._resident.add(uid)
        self._live_tokens.setdefault(uid, []).append(tokens.reshape(-1).cpu())
        self._live_latents.setdefault(uid, []).append(("lat", lat[0]))
        return logits

    def decode(self, uids: List[int], max_new_tokens: int, first_logits: torch.Tensor, do_sample: bool = False,
               temperature: float = 1.0, top_k: int = 0, top_p: float = 1.0,
               stop: Optional[Callable[[int, int], bool]] = None) -> List[List[int]]:
        """Batched decode of resident sessions. ``first_logits``: ``[len(uids), vocab]`` from their prefills."""
        out: List[List[int]] = [[] for _ in uids]
        logits = first_logits
        alive = list(range(len(uids)))
        for step in range(max_new_tokens):
            nxt = sample_next(logits, do_sample, temperature, top_k, top_p).cpu()
            keep = []
            for j, i in enumerate(alive):
                tok = int(nxt[j])
                out[i].append(tok)
                if not (stop and stop(uids[i], tok)):
                    keep.append((i, tok))
            if not keep or step == max_new_tokens - 1:
                break
            alive = [i for i, _ in keep]
            toks = [torch.tensor([t], dtype=torch.int32) for _, t in keep]
            logits, lats = self.engine.put([uids[i] for i in alive], toks, capture_latents=True)
            for (i, _), tk, lt in zip(keep, toks, lats):
                self._live_tokens[uids[i]].append(tk)
                self._live_latents[uids[i]].append(("lat", lt))
        return out

    def end_turn(self, uid: int, keep_resident: bool = False) -> SessionState:
        """Save what the plan needs and (unless ``keep_resident``) release the session's KV blocks."""
        eng = self.engine
        spec = eng.model().spec
        tokens = torch.cat([t.reshape(-1).to(torch.int32) for t in self._live_tokens[uid]])
        plan = self.plan_for(int(tokens.numel()), from_nvme=False)
        hidden_layers, kv_layers = plan.layers_of("hidden"), plan.layers_of("kv")
        hidden = None
        if hidden_layers:
            parts = []
            for kind, obj in self._live_latents[uid]:
                if kind == "lat":
                    lt = obj.wait() if hasattr(obj, "wait") else obj
                    parts.append(lt[hidden_layers])
                else:  # a restored state: its hidden layers cover a subset; missing ones force a full capture next time
                    prev: SessionState = obj
                    have = prev.layers_of("hidden")
                    if not set(hidden_layers) <= set(have):
                        # the previous plan did not keep these layers: fall back to what it kept
                        hidden_layers = [l for l in hidden_layers if l in have]
                        kv_layers = sorted(set(range(spec.n_layers)) - set(hidden_layers) - set(plan.layers_of("none")))
                        parts = [p[[i for i, l in enumerate(plan.layers_of("hidden")) if l in hidden_layers]] for p in parts]
                    idx = [have.index(l) for l in hidden_layers]
                    parts.append(prev.hidden[idx])
            from ..utils.memory import pinned_empty
            total = sum(p.shape[1] for p in parts)
            hidden = pinned_empty((len(hidden_layers), total, spec.hidden), parts[0].dtype)
            off = 0
            for p in parts:
                hidden[:, off:off + p.shape[1]].copy_(p)
                off += p.shape[1]
        kv = None
        if kv_layers:
            eng.model().finish_pending_restore()
            seq = eng.state_manager.get_sequence(uid)
            kv = eng.state_manager.kv_cache.offload_layers(seq.all_block_ids, kv_layers, layer_major=True)
        kinds = ["none"] * spec.n_layers
        for l in hidden_layers:
            kinds[l] = "hidden"
        for l in kv_layers:
            kinds[l] = "kv"
        st = SessionState(tokens=tokens, kinds=kinds, hidden=hidden, kv=kv, meta={"turns": len(self._live_tokens[uid])},
                          kv_layer_major=True)
        self.store.save(uid, st)
        if not keep_resident:
            eng.flush(uid)
            self._resident.discard(uid)
            self._live_latents.pop(uid, None)
            self._live_tokens.pop(uid, None)
        else:
            self._live_tokens[uid] = [tokens]
            self._live_latents[uid] = [("state", st)]
        return st
