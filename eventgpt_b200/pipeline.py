"""Host-fed serving loop: event windows arrive in pinned host memory, tokens leave to pinned host
memory, and the copies of neighbouring batches overlap the compute of the current one.

The reference processes one sample at a time on one stream
(``inference.py:49-63``: ``process_event_data`` -> ``model.generate``); its only multi-stream
consumer runs two prefills on separate CUDA streams (``pipeline/benchmark_e2e/benchmark_e2e_wallclock.py:641-649``).
Here the same work is pipelined over three streams with two buffer slots:

    copy-in  stream : events H2D of batch i+1
    compute  stream : egpt_visual_tokens of batch i        (the caller's current stream)
    copy-out stream : tokens D2H of batch i-1
"""
import torch

from . import events as E

TOKENS, LLM = 577, 4096


class VisualTokenStream:
    """Double-buffered host -> tokens -> host loop around ``EventGPTVisualFrontEnd.visual_tokens``."""

    def __init__(self, frontend, max_events, max_windows, sensor_size, n_frames=E.N_FRAMES, depth=2):
        self.fe, self.dev = frontend, frontend.device
        self.sensor, self.n_frames, self.depth = sensor_size, n_frames, depth
        d = self.dev
        self.dx = [torch.empty(max_events, dtype=torch.uint16, device=d) for _ in range(depth)]
        self.dy = [torch.empty(max_events, dtype=torch.uint16, device=d) for _ in range(depth)]
        self.dp = [torch.empty(max_events, dtype=torch.uint8, device=d) for _ in range(depth)]
        self.doff = [torch.empty(max_windows + 1, dtype=torch.int64, device=d) for _ in range(depth)]
        self.tok = [torch.empty((max_windows, n_frames + TOKENS, LLM), dtype=torch.bfloat16, device=d) for _ in range(depth)]
        self.htok = [torch.empty((max_windows, n_frames + TOKENS, LLM), dtype=torch.bfloat16).pin_memory() for _ in range(depth)]
        # data-dependent flags of a batch (empty slice -> the reference's ValueError, coordinate out of range, ...): one
        # device word per slot, copied out next to the tokens and turned into the exception in wait() / drain()
        self.status = [torch.zeros(1, dtype=torch.int32, device=d) for _ in range(depth)]
        self.hstatus = [torch.zeros(1, dtype=torch.int32).pin_memory() for _ in range(depth)]
        self.unchecked = [False] * depth
        self.s_in, self.s_out = torch.cuda.Stream(d), torch.cuda.Stream(d)
        self.ev_h2d = [torch.cuda.Event() for _ in range(depth)]
        self.ev_done = [torch.cuda.Event() for _ in range(depth)]
        self.ev_d2h = [torch.cuda.Event() for _ in range(depth)]
        self.posted = [None] * depth
        self.count = 0
        self.dt = None                                     # timestamps, only for submit_recording()

    def submit_recording(self, hx, hy, hp, ht, first_us, interval_us, step_us, n_clips, post=None):
        """Enqueue one time-sorted recording (pinned host tensors; ``ht`` any integer dtype) cut into ``n_clips`` clips
        [first_us + c*step_us, +interval_us) ON THE DEVICE: the (start, end) pairs come from ``egpt_clip_bounds`` (two binary
        searches per clip) and ``egpt_visual_tokens_ranges`` rasterises the possibly overlapping clips straight from the one
        copy of the events -- the 50 %-overlap clips of build_my_egpt_dsec_seq.py:137-202 with step_us = interval_us // 2,
        consecutive dataset/io.py:68-87 windows with step_us = interval_us.  Same slots, copies and flags as ``submit``."""
        return self.submit(hx, hy, hp, None, post=post, ht=ht, clips=(int(first_us), int(interval_us), int(step_us), int(n_clips)))

    def submit(self, hx, hy, hp, hoff, post=None, ht=None, clips=None):
        """Enqueue one batch (pinned host tensors: x, y uint16; p uint8; window offsets int64 [B+1]).
        Returns the slot whose pinned ``htok[slot][:B]`` holds the tokens once ``wait(slot)`` returns.
        ``post(tokens, slot)``, if given, runs right after the batch's compute was enqueued (e.g. the final all_gather of a
        multi-GPU job); it may return an asynchronous work handle (``.wait()``), which is waited for before the slot's token
        buffer is written again and in ``drain``."""
        i, s = self.count, self.count % self.depth
        self.count += 1
        n, B = hx.numel(), (clips[3] if clips else hoff.numel() - 1)
        if clips and (self.dt is None or self.dt[0].dtype != ht.dtype):
            self.dt = [torch.empty(self.dx[0].numel(), dtype=ht.dtype, device=self.dev) for _ in range(self.depth)]
        compute = torch.cuda.current_stream(self.dev)
        if i >= self.depth:
            self.s_in.wait_event(self.ev_done[s])          # slot's device inputs were consumed by batch i-depth
        else:
            self.s_in.wait_stream(compute)
        with torch.cuda.stream(self.s_in):
            self.dx[s][:n].copy_(hx, non_blocking=True)
            self.dy[s][:n].copy_(hy, non_blocking=True)
            self.dp[s][:n].copy_(hp, non_blocking=True)
            if clips:
                self.dt[s][:n].copy_(ht, non_blocking=True)
            else:
                self.doff[s][:B + 1].copy_(hoff, non_blocking=True)
            self.ev_h2d[s].record(self.s_in)
        compute.wait_event(self.ev_h2d[s])
        if i >= self.depth:
            compute.wait_event(self.ev_d2h[s])             # slot's token buffer was read out
        if self.posted[s] is not None:
            self.posted[s].wait()                          # ... also by the slot's asynchronous post step
            self.posted[s] = None
        if self.unchecked[s]:
            self._check(s)                                 # nobody waited for the slot's previous batch: its flags are due now
        self.status[s].zero_()
        if clips:
            windows = E.clip_bounds(self.dt[s][:n], clips[0], clips[1], clips[2], B)
        else:
            windows = self.doff[s][:B + 1]
        self.fe.visual_tokens(self.dx[s][:n], self.dy[s][:n], self.dp[s][:n], windows, self.sensor, n_frames=self.n_frames,
                              out=self.tok[s][:B], check_flags=False, status=self.status[s], ranges=bool(clips))
        if post is not None:
            self.posted[s] = post(self.tok[s][:B], s)      # e.g. the final all_gather of a multi-GPU job
        self.ev_done[s].record(compute)
        self.s_out.wait_event(self.ev_done[s])
        with torch.cuda.stream(self.s_out):
            self.htok[s][:B].copy_(self.tok[s][:B], non_blocking=True)
            self.hstatus[s].copy_(self.status[s], non_blocking=True)
            self.ev_d2h[s].record(self.s_out)
        self.unchecked[s] = True
        return s

    def _check(self, slot):
        """Raises what the reference raises for the slot's last batch (ValueError for an empty slice, ...)."""
        self.ev_d2h[slot].synchronize()
        self.unchecked[slot] = False
        E.check_status(int(self.hstatus[slot][0]), f"VisualTokenStream slot {slot}")

    def wait(self, slot):
        self._check(slot)
        return self.htok[slot]

    def join(self):
        """Stream-level join (no host synchronisation): the caller's current stream waits for every copy-out and post step
        enqueued so far -- what a timed region records its end event after."""
        for s in range(self.depth):
            if self.posted[s] is not None:
                self.posted[s].wait()
                self.posted[s] = None
        torch.cuda.current_stream(self.dev).wait_stream(self.s_out)

    def drain(self):
        for s in range(self.depth):
            if self.posted[s] is not None:
                self.posted[s].wait()
                self.posted[s] = None
        self.s_out.synchronize()
        torch.cuda.current_stream(self.dev).synchronize()
        for s in range(self.depth):
            if self.unchecked[s]:
                self._check(s)
