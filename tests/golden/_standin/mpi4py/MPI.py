"""Single-rank stand-in for mpi4py.MPI (fixture generation only; never shipped in the product path).

Provides exactly the surface the reference touches: dtype tags, SUM, COMM_WORLD with
rank/size/allreduce/Allreduce/bcast/Bcast/Barrier/Scatterv/Gatherv, Wtime.
"""
import time
import numpy as np

DOUBLE, FLOAT, BOOL, SHORT, INT, LONG = "f8", "f4", "b1", "i2", "i4", "i8"
UNSIGNED_SHORT, UNSIGNED_INT, UNSIGNED_LONG = "u2", "u4", "u8"
SUM = "sum"


def _buf(spec):
    return spec[0] if isinstance(spec, (list, tuple)) else spec


class Intracomm:
    rank = 0
    size = 1

    def allreduce(self, x, op=SUM):
        return x

    def Allreduce(self, send, recv, op=SUM):
        np.copyto(_buf(recv), _buf(send))

    def bcast(self, x, root=0):
        return x

    def Bcast(self, buf, root=0):
        return None

    def Barrier(self):
        return None

    def Scatterv(self, send, recv, root=0):
        np.copyto(recv, np.asarray(_buf(send)).reshape(recv.shape))

    def Gatherv(self, send, recv, root=0):
        np.copyto(_buf(recv), np.asarray(send).reshape(_buf(recv).shape))


COMM_WORLD = Intracomm()


def Wtime():
    return time.time()
