#!/usr/bin/env python
"""Host IP and GPU names (reference tests/print_hostinfo.py)."""
import socket

import torch

try:
    ip = socket.gethostbyname(socket.gethostname())
except OSError:
    ip = "127.0.0.1"
print(f"host {socket.gethostname()} ip {ip}")
for i in range(torch.cuda.device_count()):
    print(f"  cuda:{i} {torch.cuda.get_device_name(i)}")
