"""Timing stand-in for torch_sparse.SparseTensor: CSR storage, `@` = torch.sparse.mm (MKL on the CPU), gradient with
respect to the dense operand through the transposed CSR (built once per operator, like torch_sparse's cached CSC).
Bench infrastructure only."""
import warnings

import torch

warnings.filterwarnings("ignore", message="Sparse CSR tensor support is in beta state")


def _csr(row, col, val, nrow, ncol):
    key = row * ncol + col
    perm = torch.argsort(key)
    row, col, val = row[perm], col[perm], val[perm]
    crow = torch.zeros(nrow + 1, dtype=torch.long, device=row.device)
    crow[1:] = torch.cumsum(torch.bincount(row, minlength=nrow), 0)
    return torch.sparse_csr_tensor(crow, col, val, size=(nrow, ncol))


class _SpMM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dense, op):
        ctx.op = op
        return torch.sparse.mm(op._a, dense)

    @staticmethod
    def backward(ctx, dout):
        return torch.sparse.mm(ctx.op.transposed(), dout.contiguous()), None


class SparseTensor:
    def __init__(self, row=None, col=None, value=None, sparse_sizes=None, **_):
        nrow, ncol = int(sparse_sizes[0]), int(sparse_sizes[1])
        self._row, self._col, self._val = row, col, value.detach()
        self._sizes = (nrow, ncol)
        self._a = _csr(row, col, self._val, nrow, ncol)
        self._at = None

    def transposed(self):
        if self._at is None:
            self._at = _csr(self._col, self._row, self._val, self._sizes[1], self._sizes[0])
        return self._at

    def size(self, dim):
        return self._sizes[dim]

    def sizes(self):
        return list(self._sizes)

    def __matmul__(self, dense):
        return _SpMM.apply(dense, self)

    matmul = __matmul__
