"""A model check of the shared-memory ring of box_bulk_kernel (src/cuda/boxcopy.cu), the
experimental cp.async.bulk copy kernel that has not run on hardware yet.

The issuing thread's control flow is transcribed below (prologue of up to NST loads, then per
segment: wait for the stage's mbarrier phase, issue the store, commit, and refill the stage of the
PREVIOUS segment after `wait_group.read 1`; `wait_group.read 0` between items).  The hardware's
freedom is modelled adversarially: loads complete in any order at any later time, bulk groups
finish READING shared memory in commit order but as late as the waits allow.  Checked invariants:
  * a load is only issued into a stage whose previous store has finished reading it;
  * a stage is only waited for with the parity of the phase its pending load will complete;
  * every segment is stored exactly once, from the stage its load went to, after that load landed.
"""
import random

NST = 4


def run_block(items, rng):
    """items: list of segment counts (one per work item of this block); returns stored segments"""
    g = 0                                   # segments issued so far (kernel variable `g`)
    phase_done = [0] * NST                  # completed phases per stage mbarrier
    stage_holds = [None] * NST              # (item, k) whose data is in the stage (after landing)
    in_flight_loads = []                    # (stage, item, k): issued, not yet landed
    stores_reading = []                     # committed groups still reading smem, in commit order: (stage, item, k)
    loads_issued_on_stage = [0] * NST
    stored = []

    def land_some(force_stage=None):
        """the adversary lets some in-flight loads land; `force_stage`: that one must land now"""
        rng.shuffle(in_flight_loads)
        keep = []
        for (s, it, k) in in_flight_loads:
            if s == force_stage or rng.random() < 0.5:
                # landing into a stage that a store is still reading would corrupt that store
                assert all(rs != s for rs, _, _ in stores_reading), "load landed in a stage being read"
                stage_holds[s] = (it, k)
                phase_done[s] += 1
            else:
                keep.append((s, it, k))
        in_flight_loads[:] = keep

    def wait_group_read(n):
        """at most n groups may still be reading; the adversary retires no more than it must"""
        while len(stores_reading) > n:
            stores_reading.pop(0)           # groups complete in order
        # it may retire more, but never has to
        while stores_reading and rng.random() < 0.2:
            stores_reading.pop(0)

    def issue_load(it, k):
        s = (g + k) % NST
        assert all(rs != s for rs, _, _ in stores_reading), "refill of a stage whose store still reads"
        assert all(ls != s for ls, _, _ in in_flight_loads), "two loads in flight on one stage"
        loads_issued_on_stage[s] += 1
        in_flight_loads.append((s, it, k))

    for it, nseg in enumerate(items):
        for k in range(min(NST, nseg)):
            issue_load(it, k)
        for k in range(nseg):
            n = g + k
            s, parity = n % NST, (n // NST) & 1
            # the wait returns once the phase with this parity has completed: that must be the
            # phase of the load issued for (it, k), i.e. the (n // NST + 1)-th load on the stage
            assert loads_issued_on_stage[s] == n // NST + 1, "waiting for a load that was not issued"
            land_some(force_stage=s)
            assert phase_done[s] == n // NST + 1 and (phase_done[s] - 1) & 1 == parity
            assert stage_holds[s] == (it, k), "stage holds another segment's data"
            stored.append((it, k))
            stores_reading.append((s, it, k))          # store issued + commit_group
            if k >= 1 and k - 1 + NST < nseg:
                wait_group_read(1)
                issue_load(it, k - 1 + NST)
            land_some()
        wait_group_read(0)
        assert not in_flight_loads, "loads left over at the end of an item"
        g += nseg
    return stored


def test_ring_protocol_holds_for_random_items_and_adversarial_completion_orders():
    rng = random.Random(1234)
    for trial in range(400):
        items = [rng.choice([0, 1, 2, 3, 4, 5, 7, 8, 9, 16, 33]) for _ in range(rng.randint(1, 6))]
        stored = run_block(items, rng)
        want = [(it, k) for it, n in enumerate(items) for k in range(n)]
        assert stored == want
