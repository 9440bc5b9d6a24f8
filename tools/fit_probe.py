import torch,time,sys
sys.path.insert(0,"deep-learning-air-hockey-robot_b200")
from gym_air_hockey_b200.rl import TrainableNetwork
x=torch.randint(0,256,(4096,9,128,128),dtype=torch.uint8,device="cuda")
def t(f,n=5):
    f(); f(); torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter()-t0)/n*1e3
for name, cl, ac in (("fp32", False, False), ("fp32 channels_last", True, False), ("bf16 autocast", False, True), ("bf16 autocast channels_last", True, True)):
    m=TrainableNetwork().cuda().train()
    if cl: m=m.to(memory_format=torch.channels_last)
    def fb():
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=ac):
            xx = ((x.float()-128)/128)
            if cl: xx = xx.contiguous(memory_format=torch.channels_last)
            y=m.out(torch.relu(m.dense2(torch.relu(m.dense1(m.bn3(torch.relu(m.conv3(m.bn2(torch.relu(m.conv2(m.bn1(torch.relu(m.conv1(xx))))))))).flatten(1))))))
        y.float().backward(torch.ones_like(y, dtype=torch.float32))
    print(name, "fwd+bwd ms", round(t(fb),2))
