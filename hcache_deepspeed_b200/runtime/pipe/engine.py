"""PipelineEngine: executes PipeSchedule instruction streams (1F1B training, wavefront inference) over the stages of
a :class:`PipelineModule`, with data parallelism across replicas of each stage.

``train_batch(data_iter)`` consumes ``gradient_accumulation_steps`` micro-batches and returns the mean loss
(broadcast from the last stage); ``eval_batch`` runs the forward-only schedule. Activations cross stages through
``p2p.send / recv``; their shapes/dtypes travel once as a pickled meta message and are then assumed static.
Reference: ``deepspeed/runtime/pipe/engine.py`` (PipelineEngine :60, train_batch :330, eval_batch :418,
_exec_* :640-1290, _INSTRUCTION_MAP :1379)."""
from __future__ import annotations

from typing import Iterable, Optional

import os

import torch

from ... import comm as dist
from ...utils.logging import log_dist
from ..engine import DeepSpeedEngine
from ..dataloader import RepeatingLoader
from . import p2p, schedule
from .module import PipelineError, PipelineModule


def _to_tuple(x):
    return x if isinstance(x, tuple) else (x, )


class PipelineEngine(DeepSpeedEngine):

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        assert isinstance(self.module, PipelineModule), "model must base PipelineModule"
        assert self.zero_optimization_stage() < 2, "ZeRO-2 and ZeRO-3 are incompatible with pipeline parallelism"
        self.enable_backward_allreduce = False  # the schedule issues ReduceGrads
        self.grid = self.module._grid
        self.num_stages = self.module.num_pipeline_stages()
        self.stage_id = self.module.stage_id
        self.prev_stage, self.next_stage = self.stage_id - 1, self.stage_id + 1
        self.micro_batches = self.gradient_accumulation_steps()
        self.micro_batch_size = self.train_micro_batch_size_per_gpu()
        self.data_iterator = None
        self.batch_fn = None
        self.loss_model = self.module.loss_fn
        self.has_bool_tensors = False
        self.total_loss = None
        self.agg_additional_losses = None
        self._fwd_additional = {}
        self.agg_loss = torch.tensor(0.0, device=self.device)
        self.pipe_buffers = {"inputs": [], "labels": [], "outputs": [], "grads_in": []}
        self.num_pipe_buffers = 0
        self._recv_meta = {"act": None, "grad": None}
        self._sent_meta = {"act": False, "grad": False}
        if self.num_stages > 1:
            p2p.init_process_groups(self.grid)
        if self.training_data is not None:
            self._build_data_iter(self.training_data)
        log_dist(f"pipeline stage {self.stage_id}/{self.num_stages}: layers [{self.module._local_start}, "
                 f"{self.module._local_stop}) micro_batches={self.micro_batches}", ranks=[self.global_rank])

    # ------------------------------------------------------------------------------------------------
    def is_first_stage(self):
        return self.stage_id == 0

    def is_last_stage(self):
        return self.stage_id == self.num_stages - 1

    def set_dataiterator(self, iterator):
        if self.is_first_stage() or self.is_last_stage():
            self.training_dataloader = iterator
            self.data_iterator = iterator

    def set_batch_fn(self, fn):
        self.batch_fn = fn

    def set_dataloader(self, loader):
        """Only the first and the last stage read data (inputs there, labels here)."""
        if self.is_first_stage() or self.is_last_stage():
            self.training_dataloader = loader
            self.data_iterator = iter(loader)

    def set_has_attention_mask(self, value: bool):
        assert isinstance(value, bool)
        self.has_attention_mask = value

    def set_train_batch_size(self, train_batch_size):
        """Change the global batch by changing the number of micro-batches; the micro-batch size stays."""
        super().set_train_batch_size(train_batch_size)
        self.micro_batches = self.gradient_accumulation_steps()

    def reset_activation_shape(self):
        """Forget the tensor shapes exchanged between stages: call when the activation shape is about to change (e.g. a
        sequence-length curriculum). The next send / receive re-negotiates them (ref: pipe/engine.py:322)."""
        self._recv_meta = {"act": None, "grad": None}
        self._sent_meta = {"act": False, "grad": False}
        for k in self.pipe_buffers:
            self.pipe_buffers[k] = [None] * self.num_pipe_buffers

    def get_additional_losses(self):
        """Extra named losses the last stage's loss function returned next to the main one, averaged over the batch."""
        return getattr(self, "agg_additional_losses", None)

    def log_for_device(self, *msg):
        """Debug print restricted by ``HDS_PIPE_LOG_STAGE`` / ``HDS_PIPE_LOG_DP`` (-1 = every stage / replica)."""
        stage = int(os.environ.get("HDS_PIPE_LOG_STAGE", "-1"))
        dp = int(os.environ.get("HDS_PIPE_LOG_DP", "-1"))
        if stage in (-1, self.stage_id) and dp in (-1, self.grid.data_parallel_id):
            print(f"RANK={self.global_rank} PIPE-ID={self.stage_id} DATA-ID={self.grid.data_parallel_id} "
                  f"STEP-ID={self.global_steps} ::", *msg, flush=True)

    def tput_log(self, *msg):
        if self.global_rank == 0 and self.global_steps % max(1, self.steps_per_print() or 1) == 0:
            print(*msg)

    def mem_status(self, msg, print_rank=-1, reset_max=False):
        """Allocator numbers of this rank, printed on ``print_rank`` (or everywhere for -1); returns them."""
        if not torch.cuda.is_available():
            return None
        if reset_max:
            torch.cuda.reset_peak_memory_stats()
        gb = 1024 ** 3
        stats = {"allocated_gb": torch.cuda.memory_allocated() / gb, "max_allocated_gb": torch.cuda.max_memory_allocated() / gb,
                 "reserved_gb": torch.cuda.memory_reserved() / gb, "max_reserved_gb": torch.cuda.max_memory_reserved() / gb}
        if print_rank in (-1, self.global_rank):
            print(f"RANK={self.global_rank} STAGE={self.stage_id} STEP={self.global_steps} MEMSTATS {msg} "
                  + " ".join(f"{k}={v:.4f}" for k, v in stats.items()), flush=True)
        return stats

    def load_module_state_dict(self, checkpoint, strict=True, custom_load_fn=None, fetch_z3_params=False):
        """Pipeline weights live one file per layer: a checkpoint whose ``module`` entry is empty (or a directory path)
        is loaded with ``PipelineModule.load_state_dir``; a plain state dict falls through to the base engine."""
        assert custom_load_fn is None, "custom_load_fn not supported w. pipeline parallelism"
        sd = checkpoint.get("module") if isinstance(checkpoint, dict) else checkpoint
        if sd is not None and not isinstance(sd, str):
            return super().load_module_state_dict(checkpoint, strict=strict)
        load_dir = sd if isinstance(sd, str) else getattr(self, "_curr_ckpt_path", None)
        if load_dir is None:
            raise ValueError("no per-layer checkpoint directory: pass its path as checkpoint['module']")
        self.module.load_state_dir(load_dir, self.checkpoint_engine, strict=strict)

    def set_has_bool_tensors(self, v: bool):
        self.has_bool_tensors = v

    def _build_data_iter(self, dataset):
        loader = self.deepspeed_io(dataset)
        self.set_dataiterator(iter(RepeatingLoader(loader)))

    def is_gradient_accumulation_boundary(self):
        return getattr(self, "_force_grad_boundary", False)

    def _reserve_pipe_buffers(self, n: int):
        if self.num_pipe_buffers >= n:
            return
        add = n - self.num_pipe_buffers
        for k in self.pipe_buffers:
            self.pipe_buffers[k].extend([None] * add)
        self.num_pipe_buffers = n

    # ------------------------------------------------------------------------------------------------
    def train_batch(self, data_iter: Optional[Iterable] = None):
        if not torch._C.is_grad_enabled():
            raise RuntimeError("train_batch() requires gradients enabled. Use eval_batch() instead.")
        if data_iter is not None:
            self.set_dataiterator(data_iter)
        self.module.train()
        self.total_loss = None
        self._fwd_losses = []
        self._fwd_additional = {}
        sched = schedule.TrainSchedule(micro_batches=self.micro_batches, stages=self.num_stages, stage_id=self.stage_id)
        self._exec_schedule(sched)
        self.agg_train_loss = self._aggregate_total_loss()
        if self.global_steps % max(1, self.steps_per_print() or 1) == 0:
            log_dist(f"steps: {self.global_steps} loss: {float(self.agg_train_loss):0.4f}", ranks=[0])
        return self.agg_train_loss

    def eval_batch(self, data_iter, return_logits: bool = False, compute_loss: bool = True, reduce_output: str = "avg",
                   bcast_loss: bool = True, num_micro_batches: Optional[int] = None):
        self.module.eval()
        train_iter = self.data_iterator
        self.set_dataiterator(data_iter)
        mbs = self.micro_batches if num_micro_batches is None else num_micro_batches
        sched = schedule.InferenceSchedule(micro_batches=mbs, stages=self.num_stages, stage_id=self.stage_id)
        self._fwd_losses, self._eval_outputs = [], []
        self._compute_loss = compute_loss
        with torch.no_grad():
            self._exec_schedule(sched)
        out = None
        if self.is_last_stage():
            vals = self._fwd_losses if compute_loss else self._eval_outputs
            if vals and reduce_output == "avg" and all(torch.is_tensor(v) and v.dim() == 0 for v in vals):
                out = torch.stack([v.float() for v in vals]).mean()
                if self.grid is not None and self.grid.data_parallel_size > 1:
                    dist.all_reduce(out, group=self.grid.get_data_parallel_group())
                    out /= self.grid.data_parallel_size
            else:
                out = vals
        if compute_loss and bcast_loss and reduce_output == "avg":
            out = self._bcast_pipe_scalar(out if torch.is_tensor(out) else torch.tensor(0.0, device=self.device))
        self._compute_loss = True
        self.set_dataiterator(train_iter)
        if return_logits:
            return out, (self._eval_outputs if self.is_last_stage() else None)
        return out

    def _bcast_pipe_scalar(self, data: torch.Tensor, src_stage: Optional[int] = None):
        if self.num_stages == 1:
            return data
        src_stage = self.num_stages - 1 if src_stage is None else src_stage
        src = self.grid.stage_to_global(src_stage)
        t = data.detach().clone().float().to(self.device) if self.stage_id == src_stage else torch.zeros((), device=self.device)
        dist.broadcast(t, src=src, group=self.grid.get_pipe_parallel_group())
        return t

    def _aggregate_total_loss(self):
        if self.is_last_stage():
            loss = torch.stack([l.detach().float() for l in self._fwd_losses]).mean() if self._fwd_losses else \
                torch.tensor(0.0, device=self.device)
            if self.grid is not None and self.grid.data_parallel_size > 1:
                dist.all_reduce(loss, group=self.grid.get_data_parallel_group())
                loss /= self.grid.data_parallel_size
        else:
            loss = torch.tensor(0.0, device=self.device)
        self._aggregate_additional_losses()
        return self._bcast_pipe_scalar(loss)

    def _aggregate_additional_losses(self):
        """Mean over micro-batches and data-parallel replicas of every named side loss, then last stage -> all stages.
        As in the reference, ``module.get_additional_losses()`` must name the same losses on every stage."""
        from collections import OrderedDict
        if self.is_last_stage():
            names = list(self._fwd_additional)
        else:
            extra = self.module.get_additional_losses()
            names = list(extra) if extra is not None else []
        if not names:
            self.agg_additional_losses = None
            return
        if self.is_last_stage():
            vals = torch.stack([torch.stack(self._fwd_additional[n]).mean() for n in names]).to(self.device)
            if self.grid is not None and self.grid.data_parallel_size > 1:
                dist.all_reduce(vals, group=self.grid.get_data_parallel_group())
                vals /= self.grid.data_parallel_size
        else:
            vals = torch.zeros(len(names), device=self.device)
        if self.num_stages > 1:
            dist.broadcast(vals, src=self.grid.stage_to_global(self.num_stages - 1), group=self.grid.get_pipe_parallel_group())
        self.agg_additional_losses = OrderedDict((n, vals[i].clone()) for i, n in enumerate(names))

    # ------------------------------------------------------------------------------------------------
    # instruction handlers
    # ------------------------------------------------------------------------------------------------
    def _next_batch(self):
        batch = next(self.data_iterator)
        if self.batch_fn is not None:
            batch = self.batch_fn(batch)
        return batch

    def _to_device(self, x):
        if torch.is_tensor(x):
            return x.to(self.device)
        if isinstance(x, (tuple, list)):
            return type(x)(self._to_device(v) for v in x)
        return x

    def _exec_load_micro_batch(self, buffer_id):
        batch = self._next_batch()
        if self.is_first_stage():
            inputs = self._to_device(batch[0])
            if torch.is_tensor(inputs) and inputs.is_floating_point():
                inputs = inputs.clone().detach().requires_grad_(False)
            self.pipe_buffers["inputs"][buffer_id] = inputs
        if self.is_last_stage():
            self.pipe_buffers["labels"][buffer_id] = self._to_device(batch[1])

    def _exec_forward_pass(self, buffer_id):
        inputs = self.pipe_buffers["inputs"][buffer_id]
        outputs = DeepSpeedEngine.forward(self, inputs)
        self.pipe_buffers["outputs"][buffer_id] = outputs
        if self.is_last_stage():
            if getattr(self, "_compute_loss", True) and self.loss_model is not None:
                loss = self.loss_model(outputs, self.pipe_buffers["labels"][buffer_id])
                self.pipe_buffers["outputs"][buffer_id] = loss
                self._fwd_losses.append(loss.detach())
                extra = self.module.get_additional_losses()
                if extra is not None:   # named side losses of this micro-batch (ref: pipe/engine.py:796-802)
                    for name, v in extra.items():
                        self._fwd_additional.setdefault(name, []).append(v.detach().float())
            else:
                if hasattr(self, "_eval_outputs"):
                    self._eval_outputs.append(outputs.detach() if torch.is_tensor(outputs) else outputs)
                if torch.is_tensor(outputs) and outputs.dim() == 0:
                    self._fwd_losses.append(outputs.detach())

    def _exec_backward_pass(self, buffer_id):
        outputs = self.pipe_buffers["outputs"][buffer_id]
        if self.is_last_stage():
            loss = outputs / self.micro_batches
            if self._uses_wrapper_backward():
                self.optimizer.backward(loss)
            else:
                loss.backward()
        else:
            grads = self.pipe_buffers["grads_in"][buffer_id]
            outs = [o for o in _to_tuple(outputs) if torch.is_tensor(o) and o.is_floating_point() and o.requires_grad]
            gs = [g for g in _to_tuple(grads)][:len(outs)]
            scale = getattr(self.optimizer, "cur_scale", 1.0) if False else 1.0
            torch.autograd.backward(tensors=outs, grad_tensors=[g * scale for g in gs])
        self.pipe_buffers["outputs"][buffer_id] = None
        self.pipe_buffers["grads_in"][buffer_id] = None

    def _send_meta(self, tensors, dest_stage, kind):
        if self._sent_meta[kind]:
            return
        meta = [(tuple(t.shape), str(t.dtype).replace("torch.", ""), bool(t.requires_grad)) for t in tensors]
        p2p.send_obj(meta, dest_stage, device=self.device)
        self._sent_meta[kind] = True

    def _recv_meta_from(self, src_stage, kind):
        if self._recv_meta[kind] is None:
            self._recv_meta[kind] = p2p.recv_obj(src_stage, device=self.device)
        return self._recv_meta[kind]

    def _exec_send_activations(self, buffer_id):
        outs = [o for o in _to_tuple(self.pipe_buffers["outputs"][buffer_id]) if torch.is_tensor(o)]
        self._send_meta(outs, self.next_stage, "act")
        for o in outs:
            p2p.send(o.detach().contiguous(), self.next_stage)

    def _exec_recv_activations(self, buffer_id):
        meta = self._recv_meta_from(self.prev_stage, "act")
        tensors = []
        for shape, dt, rg in meta:
            t = torch.empty(shape, dtype=getattr(torch, dt), device=self.device)
            p2p.recv(t, self.prev_stage)
            if t.is_floating_point():
                t.requires_grad_(True)
            tensors.append(t)
        self.pipe_buffers["inputs"][buffer_id] = tensors[0] if len(tensors) == 1 else tuple(tensors)

    def _exec_send_grads(self, buffer_id):
        inputs = [t for t in _to_tuple(self.pipe_buffers["inputs"][buffer_id]) if torch.is_tensor(t) and t.is_floating_point()]
        grads = [t.grad if t.grad is not None else torch.zeros_like(t) for t in inputs]
        self._send_meta(grads, self.prev_stage, "grad")
        for g in grads:
            p2p.send(g.contiguous(), self.prev_stage)
        self.pipe_buffers["inputs"][buffer_id] = None

    def _exec_recv_grads(self, buffer_id):
        meta = self._recv_meta_from(self.next_stage, "grad")
        grads = []
        for shape, dt, _ in meta:
            g = torch.empty(shape, dtype=getattr(torch, dt), device=self.device)
            p2p.recv(g, self.next_stage)
            grads.append(g)
        self.pipe_buffers["grads_in"][buffer_id] = tuple(grads)

    def _exec_reduce_tied_grads(self):
        self.module.allreduce_tied_weight_gradients()

    def _exec_reduce_grads(self):
        self._force_grad_boundary = True
        from ..zero.stage_1_and_2 import DeepSpeedZeroOptimizer
        if isinstance(self.optimizer, DeepSpeedZeroOptimizer):
            pass  # stage 1 reduces inside step()
        elif dist.is_initialized() and self.dp_world_size > 1:
            self.buffered_allreduce_fallback()
        self._force_grad_boundary = False

    def _exec_optimizer_step(self, lr_kwargs=None):
        self._force_grad_boundary = True
        self._take_model_step(lr_kwargs)
        self.micro_steps += self.micro_batches
        self._force_grad_boundary = False

    _INSTRUCTION_MAP = {
        schedule.OptimizerStep: _exec_optimizer_step,
        schedule.ReduceGrads: _exec_reduce_grads,
        schedule.ReduceTiedGrads: _exec_reduce_tied_grads,
        schedule.LoadMicroBatch: _exec_load_micro_batch,
        schedule.ForwardPass: _exec_forward_pass,
        schedule.BackwardPass: _exec_backward_pass,
        schedule.SendActivation: _exec_send_activations,
        schedule.RecvActivation: _exec_recv_activations,
        schedule.SendGrad: _exec_send_grads,
        schedule.RecvGrad: _exec_recv_grads,
    }

    def _exec_schedule(self, pipe_schedule):
        self._reserve_pipe_buffers(pipe_schedule.num_pipe_buffers())
        for step_cmds in pipe_schedule:
            for cmd in step_cmds:
                fn = self._INSTRUCTION_MAP.get(type(cmd))
                if fn is None:
                    raise RuntimeError(f"{self.__class__.__name__} does not understand instruction {cmd!r}")
                fn(self, **cmd.kwargs)

    # the per-micro-batch API of the base engine makes no sense for a pipeline
    def forward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def backward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def step(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def module_state_dict(self, exclude_frozen_parameters=False):
        """The pipeline's weights are written one file per layer (``PipelineModule.save_state_dict``), so a checkpoint can
        be reloaded under another number of stages; the per-rank model-states file carries no module tensors."""
        ckpt_dir = getattr(self, "_curr_ckpt_path", None)
        if ckpt_dir is None:
            return self.module.state_dict()
        self.module.save_state_dict(ckpt_dir, self.checkpoint_engine, exclude_frozen_params=exclude_frozen_parameters)
        return None

    def save_checkpoint(self, save_dir, tag=None, client_state=None, save_latest=True, exclude_frozen_parameters=False):
        if tag is None:
            tag = f"global_step{self.global_steps}"
        self._curr_ckpt_path = os.path.join(save_dir, str(tag))
        try:
            return super().save_checkpoint(save_dir, tag=tag, client_state=client_state, save_latest=save_latest,
                                           exclude_frozen_parameters=exclude_frozen_parameters)
        finally:
            self._curr_ckpt_path = None


def is_even(number: int) -> bool:
    return number % 2 == 0
