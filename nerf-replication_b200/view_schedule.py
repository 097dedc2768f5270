"""Which view(s) a rank trains on at each iteration.

The reference gets this from a DataLoader driven by a sampler stack (src/datasets/make_dataset.py:37-71 building
src/datasets/samplers.py:50-131 on torch's samplers).  Here the images never leave HBM, so all that is left of that
stack is its stream of integers; ``ViewSchedule`` produces that stream with no DataLoader, workers or collate behind it.

Exactness: the distributed path (``DistributedSampler`` re-seeded per epoch, train.py:60-61) and one pass of the
iteration-based sampler are bit-equal to the reference's own sampler objects (tests/golden/loop_samplers.json).  The
single-process shuffled path has the same distribution but NOT the same stream from the second epoch on: the reference's
``IterationBasedBatchSampler`` loops ``while iteration <= num_iterations`` (samplers.py:60-72) and therefore starts one
extra ``RandomSampler`` pass whenever ``ep_iter`` is a multiple of the pass length (lego.yaml: 500 = 5 x 100), and
``DataLoader.__iter__`` draws a base seed per epoch -- both consume the global RNG, which this class does not emulate.
"""
import torch


def sequential_pass(n_items):
    return list(range(n_items))


def random_pass(n_items, generator=None):
    """One pass of torch.utils.data.RandomSampler: a randperm from `generator`, or -- as torch does when none is given --
    from a fresh generator seeded off the global RNG."""
    if generator is None:
        generator = torch.Generator()
        generator.manual_seed(int(torch.empty((), dtype=torch.int64).random_().item()))
    return torch.randperm(n_items, generator=generator).tolist()


def distributed_pass(n_items, epoch, rank, world, shuffle=True):
    """One pass of the reference's DistributedSampler (samplers.py:105-125): permutation seeded by the epoch, wrapped
    around to a multiple of the world size, then the rank's contiguous share."""
    if shuffle:
        order = torch.randperm(n_items, generator=torch.Generator().manual_seed(epoch)).tolist()
    else:
        order = sequential_pass(n_items)
    share = -(-n_items // world)
    order = order + order[: share * world - n_items]
    return order[share * rank: share * (rank + 1)]


class ViewSchedule:
    """Iterable of index batches.  max_iter == -1: one pass (the val/test loader); otherwise passes are repeated until
    exactly max_iter batches were handed out (IterationBasedBatchSampler, samplers.py:50-72; train.py passes
    cfg.ep_iter).  A trailing partial batch of a pass is kept unless drop_last."""

    def __init__(self, n_items, batch_size=1, shuffle=True, is_distributed=False, max_iter=-1, rank=0, world=1,
                 drop_last=False, generator=None):
        self.n_items, self.batch_size, self.shuffle, self.drop_last = n_items, batch_size, shuffle, drop_last
        self.is_distributed, self.rank, self.world = is_distributed, rank, world
        self.max_iter, self.generator, self.epoch = max_iter, generator, 0

    def set_epoch(self, epoch):                 # train.py:58-59
        self.epoch = epoch

    def _pass(self):
        if self.is_distributed:
            return distributed_pass(self.n_items, self.epoch, self.rank, self.world, self.shuffle)
        return random_pass(self.n_items, self.generator) if self.shuffle else sequential_pass(self.n_items)

    def _batches_of_pass(self):
        order = self._pass()
        for k in range(0, len(order), self.batch_size):
            chunk = order[k: k + self.batch_size]
            if len(chunk) == self.batch_size or not self.drop_last:
                yield chunk

    def __iter__(self):
        if self.max_iter == -1:
            yield from self._batches_of_pass()
            return
        produced = 0
        while produced < self.max_iter:
            before = produced
            for chunk in self._batches_of_pass():
                yield chunk
                produced += 1
                if produced == self.max_iter:
                    return
            if produced == before:
                raise RuntimeError("ViewSchedule: a pass produced no batch (empty dataset?)")

    def __len__(self):
        if self.max_iter != -1:
            return self.max_iter
        n = -(-self.n_items // self.world) if self.is_distributed else self.n_items
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)
