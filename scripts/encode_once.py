import os, sys
ROOT = "/root/repo"
sys.path[:0] = [ROOT, os.path.join(ROOT, "diffusion-model_b200")]
import torch
from models.unet_cond_geom import UnetCondWithGeomHead
from models.vae import VAE
dev = torch.device("cuda", 0)
torch.manual_seed(0); UnetCondWithGeomHead(); vae = VAE().eval().to(dev)
x = torch.rand((32, 3, 224, 224), device=dev); eps = torch.randn((32, 4, 28, 28), device=dev)
for _ in range(2):
    z, kl = vae.encode(x, eps)
torch.cuda.synchronize()
