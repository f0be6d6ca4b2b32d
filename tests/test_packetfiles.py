"""Packet files written from / read into the resident ensemble: artis_gpu_packets_write_temp / _read_temp /
_write_text against the formats of the reference -- write_temp_packetsfile (sn3d.cc:387-398: fwrite of the 304-byte
records) and write_packets (packet.cc:152-196: header + 45 columns with %d / %g / %lg conversions, re-stated here in
Python, whose % operator implements the C conversions)."""
import os

import numpy as np
import pytest

import goldenlib
from artis_b200 import abi
from artis_b200.transport import Transport

HEADER = ("#number where type_id posx posy posz dirx diry dirz last_cross tdecay e_cmf e_rf nu_cmf nu_rf "
          "escape_type_id escape_time scat_count next_trans interactions last_event emissiontype trueemissiontype "
          "em_posx em_posy em_posz absorption_type absorption_freq nscatterings em_time absorptiondirx absorptiondiry "
          "absorptiondirz stokes1 stokes2 stokes3 pol_dirx pol_diry pol_dirz originated_from_positron "
          "true_emission_velocity trueem_time pellet_nucindex\n")


def text_row(p):
    """one line of write_packets (packet.cc:163-194)"""
    g = lambda x: "%g " % float(x)  # noqa: E731
    d = lambda x: "%d " % int(x)  # noqa: E731
    s = d(p["number"]) + d(p["where"]) + d(p["type"])
    s += "".join(g(x) for x in p["pos"]) + "".join(g(x) for x in p["dir"])
    s += d(p["last_cross"]) + g(p["tdecay"]) + g(p["e_cmf"]) + g(p["e_rf"]) + g(p["nu_cmf"]) + g(p["nu_rf"])
    s += d(p["escape_type"]) + d(p["escape_time"]) + d(p["scat_count"]) + d(p["next_trans"]) + d(p["interactions"])
    s += d(p["last_event"]) + d(p["emissiontype"]) + d(p["trueemissiontype"])
    s += "".join(g(x) for x in p["em_pos"]) + d(p["absorptiontype"]) + g(p["absorptionfreq"]) + d(p["nscatterings"]) + d(p["em_time"])
    s += "".join(g(x) for x in p["absorptiondir"]) + "".join(g(x) for x in p["stokes"]) + "".join(g(x) for x in p["pol_dir"])
    s += d(p["originated_from_particlenotgamma"]) + g(p["trueemissionvelocity"]) + d(p["trueem_time"]) + d(p["pellet_nucindex"])
    return s + "\n"


def check(lib, tmp_path, n_big=None):
    fx = goldenlib.load_fixture("classic_vpkt_ts30.npz")  # packets with Stokes vectors, emission and absorption records
    pk = fx["pre"].copy() if n_big is None else np.tile(fx["pre"], -(-n_big // fx["pre"].size))[:n_big].copy()
    T = Transport(lib)
    T.init(fx["static"])
    T.packets_upload(pk)
    tmp = os.path.join(str(tmp_path), "packets_0000_ts30.tmp")
    T.packets_write_temp(tmp)
    assert open(tmp, "rb").read() == pk.tobytes()  # byte for byte what fwrite(pkt, sizeof(struct packet), npkts) writes
    T.packets_upload(pk[:1].copy())  # forget the ensemble, then take it back from the file
    T.packets_read_temp(tmp, pk.size)
    assert T.packets_download().tobytes() == pk.tobytes()
    with pytest.raises(Exception):
        T.packets_read_temp(tmp, pk.size + 1)  # a short file is an error (assert_always in the reference)
    if n_big is None:
        txt = os.path.join(str(tmp_path), "packets00_0000.out")
        T.packets_write_text(txt)
        want = HEADER + "".join(text_row(p) for p in pk)
        assert open(txt).read() == want
    T.finalize()


def test_restatement_packet_files(oracle_lib, tmp_path):
    check(oracle_lib, tmp_path)


@pytest.mark.gpu
def test_cuda_packet_files(gpu_lib, tmp_path):
    check(gpu_lib, tmp_path)
    check(gpu_lib, tmp_path, n_big=200_003)  # several host chunks, a ragged last one
