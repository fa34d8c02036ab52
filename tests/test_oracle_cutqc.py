"""Pins the cutqc oracle's stage arithmetic and the product's file writers on the reference's
OWN C programs (oracle/_ref, compiled from /root/reference/cutqc/*.c): the raw instance files
of a cut circuit go through ./measure -> ./vertical_collapse -> ./merge -> ./build as
cutqc/main.py launches them, and the result must equal the oracle pipeline (float32 + '%e'
text => 2e-6) and the uncut distribution."""
import numpy as np
import pytest

from oracle import cutqc_oracle as co
from oracle import ref_cutqc
from oracle import statevector as sv
import cutqc_cases as cc
from qdc_b200.cutqc import distributor as dist

pytestmark = pytest.mark.skipif(not ref_cutqc.available(), reason="oracle/_ref binaries not built")


@pytest.mark.parametrize("name,limit", [("ghz3", None), ("ckt5", None), ("there_and_back", None), ("chain3", None),
                                        ("dqva6_fused", None), ("chain3", 4), ("dqva6", 5)])
@pytest.mark.parametrize("et", [0, 1])
def test_reference_c_pipeline_matches_oracle(tmp_path, name, limit, et):
    circ, frags, wpm = cc.CASES[name]()
    cpm = co.complete_path_map_from_wire_path_map(wpm)
    counter, kron, terms, K, inter = co.evaluate(frags, cpm, early_termination=et)
    (schedule, want, want_dict), = co.post_process(list(circ.qubits), frags, cpm, counter, kron, terms, K,
                                                   qubit_limit=limit)
    raw = {}
    for s, sub in enumerate(frags):
        raw[s] = {}
        O_q, rho_q = co.find_subcircuit_O_rho_qubits(cpm, s)
        combos, _ = co.find_all_combinations(O_q, rho_q, list(sub.qubits))
        for inits, meas in combos:
            probs = co.instance_probabilities(sub, inits, meas)
            for mutated in co.mutate_measurement_basis(meas):
                raw[s][inter["indexed"][s][(tuple(inits), tuple(mutated))]] = (inits, mutated, probs)

    def write_layer(vc_folder, dd_folder):                 # the PRODUCT's writer, parsed by the C code
        dist.write_files(0, 1, counter, vc_folder, dd_folder, {0: schedule}, terms, K, len(frags))

    got = ref_cutqc.run_pipeline(str(tmp_path), circ.num_qubits, counter, raw, inter["kronecker_terms"],
                                 write_layer, early_termination=et)
    assert got.shape == want.shape
    assert np.abs(got - want).max() < 2e-6
    if limit is None:
        n = circ.num_qubits
        full = sv.probabilities(sv.simulate(n, circ.to_instructions()))
        rebuilt = np.zeros(1 << n)
        for k, v in co.reorder(list(circ.qubits), got, cpm, frags, schedule).items():
            rebuilt[int(k, 2)] = v
        assert np.abs(rebuilt - full).max() < 2e-6
