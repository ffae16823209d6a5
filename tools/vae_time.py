"""Times the one-frame VAE decode at full size (1024x1024: latents 128x128) on cuda:0 — the post-loop cost
of SURVEY §8 row f1 next to the 50-step denoising loop.  Random-init weights of the real architecture."""
import sys, os, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200.vae import B200QwenImageVAEDecoder

def main():
    torch.manual_seed(0)
    vae = B200QwenImageVAEDecoder(device="cuda:0")
    for p in vae.parameters():
        if p.dim() > 1 and p.shape[-1] > 1 or p.dim() == 4:
            p.normal_(0, 0.02)
    n_par = sum(p.numel() for p in vae.parameters())
    for hl in (128, 166):
        z = torch.randn(1, 16, 1, hl, hl, device="cuda:0")
        for _ in range(2):
            out = vae.decode(z).sample
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            out = vae.decode(z).sample
        e1.record()
        torch.cuda.synchronize()
        print(f"VAE_DECODE latents {hl}x{hl} -> image {tuple(out.shape)}: {e0.elapsed_time(e1) / 5:.2f} ms "
              f"(params {n_par / 1e6:.1f} M, peak mem {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB)")

if __name__ == "__main__":
    main()
