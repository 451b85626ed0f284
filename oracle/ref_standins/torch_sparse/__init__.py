"""Stand-in for torch_sparse.SparseTensor: COO sorted by (row, col), `@` = sum-SpMM with
autograd w.r.t. the dense operand, `.size(i)`.  Test infrastructure only."""
import torch


class SparseTensor:
    def __init__(self, row=None, col=None, value=None, sparse_sizes=None, **_):
        nrow, ncol = sparse_sizes
        key = row * ncol + col
        perm = torch.argsort(key, stable=True)
        self._row, self._col, self._val = row[perm], col[perm], value[perm]
        self._sizes = (int(nrow), int(ncol))

    def size(self, dim):
        return self._sizes[dim]

    def sizes(self):
        return list(self._sizes)

    def coo(self):
        return self._row, self._col, self._val

    def to_dense(self):
        out = self._val.new_zeros(self._sizes)
        out.index_put_((self._row, self._col), self._val, accumulate=True)
        return out

    def to(self, *args, **kwargs):
        res = object.__new__(SparseTensor)
        res._row, res._col = self._row, self._col
        res._val = self._val.to(*args, **kwargs)
        res._sizes = self._sizes
        return res

    def __matmul__(self, dense):
        gathered = dense[self._col] * self._val.unsqueeze(-1)
        out = dense.new_zeros((self._sizes[0],) + tuple(dense.shape[1:]))
        return out.index_add(0, self._row, gathered)

    matmul = __matmul__
