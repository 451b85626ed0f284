"""A discrete-event model of the hand-off protocol of tc_pair_gemm_kernel (deltaconv_b200/csrc/gemm_tc.cu): the mbarriers, the
three rings (raw A tiles, B halves, tensor-memory A slots), the two CTAs of a pair and every role that waits or arrives —
A producer, B producer, two converter groups per CTA, the MMA-issuing thread of the leader, the epilogues.  No arithmetic: a
slot carries the index of the k-block whose data it holds, and the model checks that

  * a consumer only ever reads a slot that holds ITS k-block (at the time it reads, and for the MMAs until they complete),
  * no arrival is made on a barrier whose pending-arrival count is already zero (the hardware fault: a second expect_tx on a
    "tile landed" barrier whose first transaction has not completed),
  * every role terminates (no deadlock).

It exists because the kernel cannot run in the GPU-less build container while its protocol can be checked there: under random
schedules and random (occasionally very long) TMA latencies, `shared_full=True` — the barrier layout before the fix, one "tile
landed" barrier per slot shared by both converter groups — is caught consuming unfilled slots and faulting, exactly what a B200
showed once per ~10^5 launches; the layout of the kernel (`shared_full=False`, a barrier per group and slot) is not.

mbarrier semantics modelled (PTX): a phase completes when the pending-arrival count and the transaction count are both zero;
the count is then re-armed; `test_wait.parity p` succeeds iff the parity of the phase currently in progress differs from p — so it
also succeeds for a waiter that is one whole phase EARLY, which is the hole the shared barriers left open."""
import random

PA_SLOTS, PB_SLOTS, PT_SLOTS, PD_SLOTS = 3, 3, 2, 4          # as in gemm_tc.cu


class ProtocolViolation(Exception):
    pass


class MBarrier:
    def __init__(self, name, count):
        self.name, self.count, self.pending, self.tx, self.phase = name, count, count, 0, 0

    def _maybe_complete(self):
        if self.pending == 0 and self.tx == 0:
            self.phase += 1
            self.pending = self.count

    def arrive(self):
        if self.pending == 0:
            raise ProtocolViolation(f'{self.name}: arrival with no pending arrival left (phase {self.phase}, tx {self.tx})')
        self.pending -= 1
        self._maybe_complete()

    def expect_tx(self, nbytes):
        self.tx += nbytes
        self.arrive()

    def complete_tx(self, nbytes):
        self.tx -= nbytes
        self._maybe_complete()

    def test(self, parity):
        return (self.phase & 1) != parity


class PairGemmModel:
    def __init__(self, units, handoffs_per_unit, shared_full, rng, slow_load_prob=0.05, data_checks=True):
        self.units, self.nkb, self.shared_full, self.rng, self.slow = units, handoffs_per_unit, shared_full, rng, slow_load_prob
        self.data_checks = data_checks        # False: keep going after a stale read, to see what the barriers do next (the GPU did)
        self.stale_reads = 0
        self.events = []                      # (due tick, sequence, callback)
        self.tick = self.seq = 0
        groups = 1 if shared_full else 2
        B = MBarrier
        self.a_full = [[[B(f'a_full[cta{c}][grp{g}][{q}]', 1) for q in range(PA_SLOTS)] for g in range(groups)] for c in range(2)]
        self.b_full = [[[B(f'b_full[cta{c}][grp{g}][{q}]', 1) for q in range(PB_SLOTS)] for g in range(groups)] for c in range(2)]
        self.a_empty = [[B(f'a_empty[cta{c}][{q}]', 1) for q in range(PA_SLOTS)] for c in range(2)]      # one agent stands for a group's 4 warps
        self.ready = [B(f'ready[{q}]', 2) for q in range(PT_SLOTS)]                                      # leader's copy: one group of each CTA
        self.done = [[B(f'done[cta{c}][{d}]', 1) for d in range(PD_SLOTS)] for c in range(2)]
        self.acc_full = [[B(f'acc_full[cta{c}][{a}]', 1) for a in range(2)] for c in range(2)]
        self.acc_empty = [B(f'acc_empty[{a}]', 2) for a in range(2)]
        self.slot_a = [[None] * PA_SLOTS for _ in range(2)]          # k-block whose raw A tile the slot holds
        self.slot_b = [[None] * PB_SLOTS for _ in range(2)]
        self.b_lo = [[None] * PB_SLOTS for _ in range(2)]
        self.tmem_a = [[None] * PT_SLOTS for _ in range(2)]
        self.mma_queue = []                   # MMAs complete in issue order

    # -- helpers ------------------------------------------------------------------------------------------
    def stale(self, message):
        self.stale_reads += 1
        if self.data_checks:
            raise ProtocolViolation(message)

    def full_bar(self, table, cta, it, slot):
        return table[cta][0 if self.shared_full else it & 1][slot]

    def full_parity(self, it, slots):
        return (it // slots) & 1 if self.shared_full else (it // (2 * slots)) & 1

    def later(self, delay, fn):
        self.seq += 1
        self.events.append((self.tick + delay, self.seq, fn))

    def load_latency(self):
        return self.rng.randint(200, 2000) if self.rng.random() < self.slow else self.rng.randint(1, 12)

    def iterations(self):
        it = 0
        for unit in range(self.units):
            for kb in range(self.nkb):
                yield unit, kb, it
                it += 1

    # -- roles: generators that yield (barrier, parity) to wait, or None to let others run ---------------------------
    def a_producer(self, c):
        for _, _, it in self.iterations():
            q = it % PA_SLOTS
            yield self.a_empty[c][q], ((it // PA_SLOTS) & 1) ^ 1
            bar = self.full_bar(self.a_full, c, it, q)
            bar.expect_tx(1)

            def land(c=c, q=q, it=it, bar=bar):
                self.slot_a[c][q] = it
                bar.complete_tx(1)
            self.later(self.load_latency(), land)
            yield None

    def b_producer(self, c):
        for _, _, it in self.iterations():
            sb = it % PB_SLOTS
            if it >= PB_SLOTS:
                prev = it - PB_SLOTS
                yield self.done[c][prev % PD_SLOTS], (prev // PD_SLOTS) & 1
            bar = self.full_bar(self.b_full, c, it, sb)
            bar.expect_tx(1)

            def land(c=c, sb=sb, it=it, bar=bar):
                self.slot_b[c][sb] = it
                bar.complete_tx(1)
            self.later(self.load_latency(), land)
            yield None

    def converter(self, c, grp):
        for _, _, it in self.iterations():
            if it & 1 != grp:
                continue
            qa, q, sb = it % PA_SLOTS, it % PT_SLOTS, it % PB_SLOTS
            yield self.full_bar(self.a_full, c, it, qa), self.full_parity(it, PA_SLOTS)
            if it >= PT_SLOTS:
                prev = it - PT_SLOTS
                yield self.done[c][prev % PD_SLOTS], (prev // PD_SLOTS) & 1
            if self.slot_a[c][qa] != it:
                self.stale(f'converter cta{c} grp{grp}: raw A slot {qa} holds k-block {self.slot_a[c][qa]}, wanted {it}')
            self.tmem_a[c][q] = it
            yield None
            self.a_empty[c][qa].arrive()
            yield self.full_bar(self.b_full, c, it, sb), self.full_parity(it, PB_SLOTS)
            if self.slot_b[c][sb] != it:
                self.stale(f'converter cta{c} grp{grp}: B slot {sb} holds k-block {self.slot_b[c][sb]}, wanted {it}')
            self.b_lo[c][sb] = it
            yield None
            self.ready[q].arrive()

    def mma(self):
        tcount = 0
        for unit, kb, it in self.iterations():
            acc = tcount & 1
            if kb == 0:
                yield self.acc_empty[acc], ((tcount >> 1) & 1) ^ 1
            q, sb = it % PT_SLOTS, it % PB_SLOTS
            yield self.ready[q], (it // PT_SLOTS) & 1
            self.check_operands(it, q, sb, 'issue')
            last = kb == self.nkb - 1
            self.mma_queue.append((it, q, sb, acc, last))
            self.later(self.rng.randint(2, 10), self.mma_complete)
            if last:
                tcount += 1
            yield None

    def check_operands(self, it, q, sb, when):
        for c in range(2):
            if not (self.tmem_a[c][q] == it and self.slot_b[c][sb] == it and self.b_lo[c][sb] == it):
                self.stale(f'MMA of k-block {it} at {when}: cta{c} holds A {self.tmem_a[c][q]}, B {self.slot_b[c][sb]}, B_lo {self.b_lo[c][sb]}')

    def mma_complete(self):
        it, q, sb, acc, last = self.mma_queue.pop(0)
        self.check_operands(it, q, sb, 'completion')            # the operands must have stayed in place while it ran
        for c in range(2):                                       # tcgen05.commit, multicast to both CTAs
            self.done[c][it % PD_SLOTS].arrive()
            if last:
                self.acc_full[c][acc].arrive()

    def epilogue(self, c):
        for tcount in range(self.units):
            acc = tcount & 1
            yield self.acc_full[c][acc], (tcount >> 1) & 1
            yield None                                            # drain the accumulator
            self.acc_empty[acc].arrive()

    # -- scheduler ------------------------------------------------------------------------------------------
    def run(self, max_ticks=2_000_000, role_weights=None):
        """role_weights: name -> relative chance of being scheduled when runnable (default 1): slow roles stress the windows in
        which a parity test is meaningful from the other side (a waiter that is LATE by two phases would block for ever)."""
        roles = {'mma': self.mma()}
        for c in range(2):
            roles[f'a_prod{c}'] = self.a_producer(c)
            roles[f'b_prod{c}'] = self.b_producer(c)
            roles[f'epi{c}'] = self.epilogue(c)
            for g in range(2):
                roles[f'conv{c}{g}'] = self.converter(c, g)
        waiting = {name: None for name in roles}                  # name -> (barrier, parity) it is blocked on
        while roles:
            self.tick += 1
            if self.tick > max_ticks:
                raise ProtocolViolation(f'no progress: {sorted(roles)} still running, blocked on ' +
                                        ', '.join(f'{n}: {w[0].name}/{w[1]}' for n, w in waiting.items() if n in roles and w))
            due = [e for e in self.events if e[0] <= self.tick]
            if due:
                self.events = [e for e in self.events if e[0] > self.tick]
                for _, _, fn in sorted(due, key=lambda e: (e[0], e[1])):
                    fn()
            runnable = [n for n in roles if waiting[n] is None or waiting[n][0].test(waiting[n][1])]
            if not runnable:
                if not self.events:
                    raise ProtocolViolation('deadlock: ' + ', '.join(f'{n}: {w[0].name}/{w[1]}' for n, w in waiting.items() if n in roles and w))
                self.tick = min(e[0] for e in self.events) - 1
                continue
            if role_weights:
                name = self.rng.choices(runnable, weights=[role_weights.get(n, 1.0) for n in runnable])[0]
            else:
                name = self.rng.choice(runnable)
            try:
                waiting[name] = next(roles[name])
            except StopIteration:
                del roles[name]
        while self.events:                                        # outstanding completions
            self.tick += 1
            due = [e for e in self.events if e[0] <= self.tick]
            self.events = [e for e in self.events if e[0] > self.tick]
            for _, _, fn in sorted(due, key=lambda e: (e[0], e[1])):
                fn()


def run_schedules(shared_full, schedules, seed=0, units=4, handoffs=4, slow_load_prob=0.05):
    """Number of schedules that violated the protocol, and the first violation's message."""
    bad, first = 0, None
    for s in range(schedules):
        rng = random.Random(seed * 1_000_003 + s)
        try:
            PairGemmModel(units, handoffs, shared_full, rng, slow_load_prob).run()
        except ProtocolViolation as exc:
            bad += 1
            first = first or f'schedule {s}: {exc}'
    return bad, first
