"""CUDA-graph replay of the similarity helpers for a FIXED input shape.

`TemporalAligner.get_text_visual_sim_joint / _dual` (tan_model.py:231-277) are what the alignment / retrieval
evaluations call per window or per video; at T <= 512 a forward is ~330 launches of a few microseconds each, i.e.
launch-bound (3.4 ms whatever the size, profiles/configs_r02.txt).  For a fixed (B, T, N) - the T-sweep of BASELINE
configs[4], equal-length windows - the two helpers are captured once and replayed: same kernels, same results, one
launch.  The model must not draw random positional offsets (`random_pos_start = 0`, the evaluation setting of the
reference: eval/eval_zeroshot_align.py builds the model with random_pos_start=0) - a host-side draw would be frozen
into the graph."""
import torch


class GraphedSimilarity:
    def __init__(self, model, B, T, N, d_video, d_text, interpolate_from=None, device="cuda"):
        if getattr(model, "random_pos_start", 0):
            raise ValueError("GraphedSimilarity needs random_pos_start = 0 (a random offset would be frozen into the graph)")
        if model.training:
            raise ValueError("GraphedSimilarity is for inference: call model.eval() first")
        self.model, self.interpolate_from = model, interpolate_from
        self.video = torch.zeros(B, T, d_video, device=device)
        self.text = torch.zeros(B, N, d_text, device=device)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):
                self._run()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.joint, self.dual = self._run()

    def _run(self):
        m = self.model
        return (m.get_text_visual_sim_joint(self.video, self.text, interpolate_from=self.interpolate_from),
                m.get_text_visual_sim_dual(self.video, self.text, interpolate_from=self.interpolate_from))

    def __call__(self, video, text):
        """Returns (joint, dual) logits [B, S, T, N]: views of the graph's static outputs, valid until the next call."""
        if tuple(video.shape) != tuple(self.video.shape) or tuple(text.shape) != tuple(self.text.shape):
            raise ValueError(f"GraphedSimilarity was captured for video {tuple(self.video.shape)} / text "
                             f"{tuple(self.text.shape)}, got {tuple(video.shape)} / {tuple(text.shape)}")
        self.video.copy_(video, non_blocking=True)
        self.text.copy_(text, non_blocking=True)
        self.graph.replay()
        return self.joint, self.dual
