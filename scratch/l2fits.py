"""Two fits of the L model at 8,192 rows (for an ncu launch list). python scratch/l2fits.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, MLP
B = 8192
torch.manual_seed(0)
net = MLP(784, [4096] * 4, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.Adam(net.parameters(), lr=1e-3), precision='bf16', device='cuda:0', max_batch=B)
x = torch.randn(B, 784, device='cuda'); y = torch.randint(0, 10, (B,), device='cuda')
for _ in range(2): p, l = model.fit(x, y)
torch.cuda.synchronize(); print('ok', l.item())
