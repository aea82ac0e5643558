"""placeholder"""
def available():
    import torch
    return torch.cuda.is_available()
def supports(*, numeric, n_classes):
    return (False, "not built yet")
