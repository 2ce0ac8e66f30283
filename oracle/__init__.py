"""CPU oracle for the porl offline-RL update step.  TEST INFRASTRUCTURE ONLY.

This package restates, in eager CPU PyTorch (the arithmetic provider the
reference itself uses, torch is un-vendored there: requirements.txt:1), the
algorithm of the reference's hot path:

  * oracle.nets      -- util/util.py:29-47 (mlp), agent/value_functions.py:31-42 (TwinV),
                        agent/policy.py:12-49 (Gaussian heads), src/porl/net/q_network.py:8-30
  * oracle.steps     -- agent/por.py:73-112, agent/sorl.py:78-152,
                        src/porl/train/cql_trainer.py:60-124, dqn_trainer.py:182-204
  * oracle.buffer    -- src/porl/buffer/replaybuffer.py:14-76, DataLoader(shuffle) index stream
  * oracle.formulas  -- closed-form gradients of the losses / Adam / cosine LR used for kernel KATs
  * oracle.costmap   -- util/costmap.py:7-64

Parity pinning: the reference holds no golden vectors for the update maths
(SURVEY.md 8c), so the oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF,
imported in the build container from /root/reference by tests/golden/make_golden.py
and committed as fixtures under tests/golden/.  tests/test_oracle_golden.py checks
the oracle against those fixtures (and live against /root/reference when it exists).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product (porl_b200/) never does: it fails loudly
when libporl_b200.so is missing.
"""
