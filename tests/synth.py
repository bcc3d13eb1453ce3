"""Seeded synthetic Bench2Drive-shaped batches shared by tests, smoke and bench."""
import torch


def make_batch(B, L=2, H=180, W=320, P=5, A=7, seed=0, smooth=True, edge_cases=True):
    g = torch.Generator().manual_seed(seed)
    if smooth:
        # low-frequency structure + noise so that bicubic / softmax see something non-trivial
        low = torch.rand(B, L, 6, 10, 3, generator=g)
        img = torch.nn.functional.interpolate(low.permute(0, 1, 4, 2, 3).reshape(B * L, 3, 6, 10),
                                              size=(H, W), mode="bilinear", align_corners=False)
        img = img.reshape(B, L, 3, H, W).permute(0, 1, 3, 4, 2)
        img = (img * 200 + torch.rand(B, L, H, W, 3, generator=g) * 55).clamp(0, 255).to(torch.uint8)
    else:
        img = torch.randint(0, 256, (B, L, H, W, 3), generator=g, dtype=torch.uint8)
    gaze = torch.rand(B, L, P, 2, generator=g)
    nvalid = torch.randint(0, P + 1, (B, L), generator=g)
    for b in range(B):
        for t in range(L):
            gaze[b, t, nvalid[b, t]:] = -1.0
    if edge_cases and B >= 4:
        gaze[0, 0] = -1.0                       # empty frame
        gaze[1, -1, 0] = torch.tensor([0.0, 0.0])
        gaze[1, -1, 1] = torch.tensor([1.0, 1.0])  # exactly on the border
        gaze[2, -1, 1] = gaze[2, -1, 0]         # duplicate point (valid or padding alike)
        gaze[2, -1, 0] = torch.tensor([0.5, 0.5]); gaze[2, -1, 1] = torch.tensor([0.5, 0.5])
        gaze[3, -1, 0] = torch.tensor([-1.0, 0.4])  # negative on one axis only => invalid
    actions = torch.rand(B, L, A, generator=g)
    return img.contiguous(), gaze.reshape(B, L, P * 2).contiguous(), actions.contiguous()
