def zeros(tensor):
    if tensor is not None:
        tensor.data.fill_(0)
