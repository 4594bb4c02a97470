"""Host-side minibatch containers with the field layout of dmlc::RowBlock<uint64_t>
(what learn/base/minibatch_iter.h hands to ProcessMinibatch) and of the localized
RowBlockContainer<unsigned> that Localizer::Localize fills (learn/base/localizer.h:42-51)."""
import numpy as np


class RowBlock:
    """size rows; offset size_t[rows+1] (offset[0] == 0); index uint64[nnz];
    value float[nnz] or None (binary features); label float[rows]."""

    def __init__(self, offset, index, value=None, label=None):
        self.offset = np.ascontiguousarray(offset, dtype=np.uint64)
        self.index = np.ascontiguousarray(index, dtype=np.uint64)
        self.value = None if value is None else np.ascontiguousarray(value, dtype=np.float32)
        self.label = None if label is None else np.ascontiguousarray(label, dtype=np.float32)
        assert self.offset.ndim == 1 and len(self.offset) >= 1 and int(self.offset[0]) == 0
        assert len(self.index) == int(self.offset[-1])
        assert self.value is None or len(self.value) == len(self.index)
        assert self.label is None or len(self.label) == self.rows

    @property
    def rows(self):
        return len(self.offset) - 1

    @property
    def nnz(self):
        return len(self.index)

    def slice(self, begin, end):
        """Rows [begin, end) as a new block (what MinibatchIter::Push assembles)."""
        o0, o1 = int(self.offset[begin]), int(self.offset[end])
        return RowBlock(self.offset[begin:end + 1] - self.offset[begin], self.index[o0:o1],
                        None if self.value is None else self.value[o0:o1],
                        None if self.label is None else self.label[begin:end])


class LocalBlock:
    """Localized minibatch: uniq uint64[U] ascending transformed keys, cnt[U], offset
    size_t[rows+1], index uint32[nnz] in 0..U-1, value or None, label."""

    def __init__(self, uniq=None, cnt=None, offset=None, index=None, value=None, label=None):
        self.uniq, self.cnt, self.offset = uniq, cnt, offset
        self.index, self.value, self.label = index, value, label

    @property
    def rows(self):
        return len(self.offset) - 1


def read_libsvm(path):
    """Tiny libsvm reader ("label idx:val ..."); drops value when all ones, as
    MinibatchIter::Next does (learn/base/minibatch_iter.h:114-116)."""
    offset, index, value, label = [0], [], [], []
    with open(path) as f:
        for line in f:
            t = line.split()
            if not t:
                continue
            label.append(float(t[0]))
            for kv in t[1:]:
                k, v = kv.split(":")
                index.append(int(k))
                value.append(float(v))
            offset.append(len(index))
    value = np.asarray(value, np.float32)
    return RowBlock(np.asarray(offset, np.uint64), np.asarray(index, np.uint64),
                    None if np.all(value == 1) else value, np.asarray(label, np.float32))
