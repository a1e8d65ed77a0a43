"""Gnss_Synchro monitor stream: the hand-written proto3 encoder against the protobuf runtime.

The schema below restates docs/protobuf/gnss_synchro.proto of the reference (field numbers, names, types); the runtime
(google.protobuf, installed) parses the product's bytes and re-serialises them: both directions must agree byte for byte,
which is what a reference-side reader (Serdes_Gnss_Synchro::readProtobuffer, the monitoring clients) relies on.
"""
import socket

import numpy as np
import pytest

from gnss_sim_receiver_b200 import _hostapi as H

pb = pytest.importorskip("google.protobuf")
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory  # noqa: E402

T = descriptor_pb2.FieldDescriptorProto
FIELDS = [("system", 1, T.TYPE_STRING), ("signal", 2, T.TYPE_STRING), ("prn", 3, T.TYPE_UINT32), ("channel_id", 4, T.TYPE_INT32),
          ("acq_delay_samples", 5, T.TYPE_DOUBLE), ("acq_doppler_hz", 6, T.TYPE_DOUBLE), ("acq_samplestamp_samples", 7, T.TYPE_UINT64),
          ("acq_doppler_step", 8, T.TYPE_UINT32), ("flag_valid_acquisition", 9, T.TYPE_BOOL), ("fs", 10, T.TYPE_INT64),
          ("prompt_i", 11, T.TYPE_DOUBLE), ("prompt_q", 12, T.TYPE_DOUBLE), ("cn0_db_hz", 13, T.TYPE_DOUBLE),
          ("carrier_doppler_hz", 14, T.TYPE_DOUBLE), ("carrier_phase_rads", 15, T.TYPE_DOUBLE), ("code_phase_samples", 16, T.TYPE_DOUBLE),
          ("tracking_sample_counter", 17, T.TYPE_UINT64), ("flag_valid_symbol_output", 18, T.TYPE_BOOL),
          ("correlation_length_ms", 19, T.TYPE_INT32), ("flag_valid_word", 20, T.TYPE_BOOL), ("tow_at_current_symbol_ms", 21, T.TYPE_UINT32),
          ("pseudorange_m", 22, T.TYPE_DOUBLE), ("rx_time", 23, T.TYPE_DOUBLE), ("flag_valid_pseudorange", 24, T.TYPE_BOOL),
          ("interp_tow_ms", 25, T.TYPE_DOUBLE), ("flag_PLL_180_deg_phase_locked", 26, T.TYPE_BOOL)]


def observables_class():
    f = descriptor_pb2.FileDescriptorProto(name="gnss_synchro.proto", package="gnss_sdr", syntax="proto3")
    m = f.message_type.add(name="GnssSynchro")
    for name, num, typ in FIELDS:
        m.field.add(name=name, number=num, type=typ, label=T.LABEL_OPTIONAL)
    o = f.message_type.add(name="Observables")
    o.field.add(name="observable", number=1, type=T.TYPE_MESSAGE, type_name=".gnss_sdr.GnssSynchro", label=T.LABEL_REPEATED)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(f)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("gnss_sdr.Observables"))


def key(name):
    return "flag_pll_180_deg_phase_locked" if name == "flag_PLL_180_deg_phase_locked" else name


def random_item(rng, ch):
    return dict(system="GE"[ch % 2], signal=["1C", "1B"][ch % 2], prn=int(rng.integers(1, 37)), channel_id=ch - 1,
                acq_delay_samples=float(rng.uniform(0, 4000)), acq_doppler_hz=float(rng.integers(-20, 20) * 250), acq_samplestamp_samples=int(rng.integers(0, 2**40)),
                acq_doppler_step=250, flag_valid_acquisition=1, fs=4000000, prompt_i=float(rng.normal(0, 900)), prompt_q=float(rng.normal(0, 90)),
                cn0_db_hz=float(rng.uniform(30, 50)), carrier_doppler_hz=float(rng.uniform(-5e3, 5e3)), carrier_phase_rads=float(rng.normal(0, 1e5)),
                code_phase_samples=float(rng.uniform(0, 1)), tracking_sample_counter=int(rng.integers(0, 2**62)), flag_valid_symbol_output=int(rng.integers(0, 2)),
                correlation_length_ms=int(rng.choice([1, 4, 20])), flag_valid_word=int(rng.integers(0, 2)), tow_at_current_symbol_ms=int(rng.integers(0, 604800000)),
                pseudorange_m=float(rng.uniform(2e7, 2.6e7)), rx_time=float(rng.uniform(0, 604800)), flag_valid_pseudorange=int(rng.integers(0, 2)),
                interp_tow_ms=float(rng.uniform(0, 6e8)), flag_pll_180_deg_phase_locked=int(rng.integers(0, 2)))


def test_wire_bytes_match_the_protobuf_runtime():
    Obs = observables_class()
    rng = np.random.default_rng(12)
    items = [random_item(rng, ch) for ch in range(6)]
    items.append(dict(system="G", signal="1C"))                       # every scalar at its default: omitted on the wire
    items.append(dict(system="E", signal="5X", channel_id=-7, correlation_length_ms=-1, fs=-5, prompt_i=-0.0))  # ten-byte varints, -0.0
    data = H.serialize_synchros(items)
    msg = Obs()
    msg.ParseFromString(data)
    assert len(msg.observable) == len(items)
    for got, want in zip(msg.observable, items):
        for name, _, typ in FIELDS:
            w = want.get(key(name), "" if typ == T.TYPE_STRING else 0)
            g = getattr(got, name)
            assert (g == bool(w)) if typ == T.TYPE_BOOL else (g == w), (name, g, w)
    assert msg.SerializeToString(deterministic=True) == data
    # one item per datagram is what the monitor sends (gnss_synchro_monitor.cc:78-82)
    one = Obs()
    one.ParseFromString(H.serialize_synchros(items[:1]))
    assert one.observable[0].prn == items[0]["prn"] and one.SerializeToString() == H.serialize_synchros(items[:1])
    assert H.serialize_synchros([]) == b""


def test_udp_sink_decimates_per_channel_and_reaches_every_client():
    Obs = observables_class()
    socks = []
    for _ in range(2):
        s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        s.bind(("127.0.0.1", 0))
        s.settimeout(2.0)
        socks.append(s)
    # the reference sends to ONE port on several addresses; two local listeners need two monitors here
    rng = np.random.default_rng(4)
    for s in socks:
        port = s.getsockname()[1]
        conf = f"TrackingMonitor.enable_monitor=true\nTrackingMonitor.client_addresses=127.0.0.1_127.0.0.1\nTrackingMonitor.udp_port={port}\nTrackingMonitor.decimation_factor=3\n"
        mon = H.SynchroMonitor(conf, "TrackingMonitor", 2)
        assert mon.enabled
        items = [random_item(rng, 1) for _ in range(7)]
        sent = mon.write(1, items[:4])   # the 3rd goes out ...
        sent = mon.write(1, items[4:])   # ... and the 6th: the count carries over between calls
        assert sent == 2 * 2             # two datagrams, two client addresses each
        got = []
        for _ in range(4):
            m = Obs()
            m.ParseFromString(s.recv(4096))
            got.append(m.observable[0].tracking_sample_counter)
        assert got == [items[2]["tracking_sample_counter"]] * 2 + [items[5]["tracking_sample_counter"]] * 2
        mon.close()
    off = H.SynchroMonitor("TrackingMonitor.udp_port=1\n", "TrackingMonitor", 1)
    assert not off.enabled and off.write(0, [random_item(rng, 0)]) == 0
    for s in socks:
        s.close()
