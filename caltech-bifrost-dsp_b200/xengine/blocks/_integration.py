"""Integration-boundary bookkeeping shared by Corr and CorrAcc.

Both reference blocks carry the same little state machine inline (corr_block.py:346-470,
corr_acc_block.py:201-333): an integration of ``acc_len`` samples starts when the sample counter equals
``start_time``; the call that starts at ``last = first + acc_len - step`` closes it; a new upstream sequence
while running re-aligns ``start_time`` to a later boundary; ``start_time = -1`` means "as soon as possible";
``acc_len = 0`` means stop.  Here it is one small class so that both blocks (and their CPU tests) share it.
"""


class IntegrationClock(object):
    """
    :param recover_skip: integrations to skip, beyond the ones already missed, when a new input sequence
        arrives mid-run: 10 for Corr (corr_block.py:367), 2 for CorrAcc (corr_acc_block.py:227).
    :param asap: how ``start_time == -1`` is resolved: 'boundary' = next multiple of ``acc_len`` after now
        (corr_block.py:397-398), 'now' = the current sample (corr_acc_block.py:245-246).
    """

    def __init__(self, recover_skip, asap):
        assert asap in ("boundary", "now")
        self.recover_skip = recover_skip
        self.asap = asap
        self.running = False
        self.start_time = 0
        self.acc_len = 0
        self.first = None
        self.last = None

    def resequence(self, now):
        """A new input sequence begins at sample ``now``.  Returns True when a running integration train was
        re-aligned (the caller then rewrites ``acc_len``/``seq0`` in its output header)."""
        if not self.running:
            return False
        missed = (now - self.start_time) // self.acc_len
        self.start_time += (missed + self.recover_skip) * self.acc_len
        self.running = False
        return True

    def retarget(self, start_time, acc_len, now):
        """New command values: integrate ``acc_len`` samples from ``start_time`` (-1 = as soon as possible)."""
        self.acc_len = acc_len
        if start_time == -1:
            if self.asap == "boundary" and acc_len > 0:
                start_time = now - (now % acc_len) + acc_len
            else:
                start_time = now
        self.start_time = start_time
        self.running = False

    def stop(self):
        self.running = False

    def begins(self, now, step):
        """True when the input that starts at ``now`` opens a new integration train; arms first/last."""
        if now != self.start_time:
            return False
        self.running = True
        self.first = now
        self.last = now + self.acc_len - step
        return True

    def advance(self, step):
        """The integration that ended at ``last`` is done: move the window to the next one."""
        self.first = self.last + step
        self.last = self.first + self.acc_len - step
