"""Diagnostic only (not product code): does a tensor-map TMA load work on this box at all, and what does a working
kernel's SASS look like?  Saves the cubin for disassembly."""
import torch, triton, triton.language as tl
from triton.tools.tensor_descriptor import TensorDescriptor

@triton.jit
def k(desc, out, BY: tl.constexpr, BX: tl.constexpr):
    t = desc.load([2, 4])
    oy = tl.arange(0, BY)[:, None]
    ox = tl.arange(0, BX)[None, :]
    tl.store(out + oy * BX + ox, t)

a = torch.arange(34 * 160, dtype=torch.float32, device="cuda").reshape(34, 160)
out = torch.empty(8 * 32, dtype=torch.float32, device="cuda")
desc = TensorDescriptor.from_tensor(a, [8, 32])
h = k[(1,)](desc, out, 8, 32)
torch.cuda.synchronize()
print("triton TMA ok:", out[0].item(), "want", a[2, 4].item(), "| out[32]", out[32].item(), "want", a[3, 4].item())
open("gpurun_out/triton_tma.cubin", "wb").write(h.asm["cubin"])
open("gpurun_out/triton_tma.ptx", "w").write(h.asm["ptx"])
