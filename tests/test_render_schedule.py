"""CPU: the arithmetic of the device-driven render loop -- the step count a round takes (render_round_kernel, csrc/render_infer.cu)
against the launch bound the host sizes its sample buffers with (Testbed::render_frame, host/testbed_render.cpp).  The kernel writes
n_alive * n_steps samples; the host only knows an upper bound of n_alive.  If the bound formula were too small the march would
write past the buffers, so the two restatements below are kept next to each other and compared over the whole parameter range."""
import random


def steps_of_round(n_pixels, n_alive, min_steps, max_steps, tail_steps, min_batch):
    """render_round_kernel: the reference's rule on max(n_pixels, min_batch), then the tail extension"""
    if n_alive == 0:
        return 0
    batch = max(n_pixels, min_batch)
    n_steps = min(max(batch // n_alive, min_steps), max_steps)
    if tail_steps > n_steps:
        n_steps = max(n_steps, min(tail_steps, batch // (4 * n_alive)))
    return n_steps


def sample_bound(n_pixels, bound, min_steps, max_steps, tail_steps, min_batch):
    """Testbed::render_frame: by_rule / by_tail"""
    batch = max(n_pixels, min_batch)
    by_rule = max(min(batch, bound * max_steps), bound * min_steps)
    by_tail = min(batch // 4, bound * tail_steps) if tail_steps > max_steps else 0
    return max(by_rule, by_tail, 1)


def test_samples_of_a_round_never_exceed_the_hosts_bound():
    g = random.Random(7)
    for _ in range(200000):
        n_pixels = g.choice([1, 7, 640, 83000, 259200, 640000, 2073600, g.randrange(1, 1 << 23)])
        min_batch = g.choice([0, 1 << 18, 1 << 21, g.randrange(0, 1 << 22)])
        min_steps = g.choice([1, 1, 1, 2, 4])
        max_steps = max(min_steps, g.choice([1, 4, 8, 8, 8, 16]))
        tail_steps = min(max(g.choice([1, 8, 32, 64, 100]), max_steps), 64)
        bound = g.randrange(1, n_pixels + 1)
        n_alive = g.randrange(0, bound + 1) if g.random() < 0.9 else bound
        n_steps = steps_of_round(n_pixels, n_alive, min_steps, max_steps, tail_steps, min_batch)
        assert n_alive * n_steps <= sample_bound(n_pixels, bound, min_steps, max_steps, tail_steps, min_batch), (n_pixels, min_batch, min_steps, max_steps, tail_steps, bound, n_alive)
        if n_alive:
            assert min_steps <= n_steps <= max(max_steps, tail_steps)


def test_the_references_rule_is_the_special_case():
    # min_batch 0 and tail_steps == max_steps: clamp(n_pixels / n_alive, MIN, MAX) of testbed.cu:547
    for n_pixels, n_alive in [(2073600, 2073600), (2073600, 431000), (2073600, 259199), (2073600, 100), (640000, 640001 // 2)]:
        assert steps_of_round(n_pixels, n_alive, 1, 8, 8, 0) == min(max(n_pixels // n_alive, 1), 8)
    # a 135-row band of a 1080p frame with most rays alive: one step per round by the reference's rule, eight with the 2^21 batch
    assert steps_of_round(259200, 200000, 1, 8, 8, 0) == 1
    assert steps_of_round(259200, 200000, 1, 8, 32, 1 << 21) == 8
