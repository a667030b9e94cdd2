"""CPU: host-side logic of the operator mirror that needs no device -- ModelId parsing / hashing, the metadata
manager rules, checkpoint snapshot / restore.  Mirrors T/models/core/ModelIdSpec.scala:26-52,
T/api/managers/MetadataManagerSpec.scala:47-77, S/api/functions/EvaluationCoFunction.scala:63-83."""
import pytest

from flink_jpmml_b200 import api

UUID = "2bc91c20-bf64-4860-ba11-de97a95d05cf"


def test_model_id_parser_matches_uuid(golden_vectors):
    g = golden_vectors["model_id"]
    assert api.ModelId.fromIdentifier(f"{g['uuid']}_{g['version']}") == api.ModelId(g["uuid"], g["version"])   # ModelIdSpec.scala:31-36
    with pytest.raises(api.WrongModelIdFormat):
        api.ModelId.fromIdentifier(g["bad"])                                                                   # ModelIdSpec.scala:38-41
    for bad in ("", UUID, UUID + "_", UUID + "_1x", UUID + "_-1", UUID.upper() + "_1", UUID[:-1] + "_1", UUID + "-1"):
        with pytest.raises(api.WrongModelIdFormat):
            api.ModelId.fromIdentifier(bad)
    assert api.ModelId.fromIdentifier(UUID + "_007").version == 7


def test_model_id_hash_is_java_string_hash():
    # ModelIdSpec.scala:46-52: modelId.hashCode == (uuid + "_" + version).hashCode
    assert api.ModelId(UUID, 23).hashCode() == api.java_string_hash(UUID + "_23")
    assert api.java_string_hash("hello") == 99162322 and api.java_string_hash("") == 0


def test_metadata_manager_rules_and_checkpoint_roundtrip():
    op = api.EvaluationCoFunction(None)
    try:
        a, b = "aaaaaaaa-0000-0000-0000-000000000001", "bbbbbbbb-0000-0000-0000-000000000002"
        op.processElement2(api.AddMessage(a, 1, "/models/a.pmml"))
        assert op.servingMetadata == 1
        op.processElement2(api.AddMessage(a, 1, "/models/other.pmml"))        # Add on a known id: ignored
        assert op.servingMetadata == 1 and b"/models/a.pmml" in op.snapshotState()
        op.processElement2(api.AddMessage(a, 2, "/models/a2.pmml"))           # a newer version is a new id
        op.processElement2(api.AddMessage(b, 1, "/models/b.pmml"))
        assert op.servingMetadata == 3
        op.processElement2(api.DelMessage(b, 7))                              # Del of an unknown id: no-op
        assert op.servingMetadata == 3
        op.processElement2(api.DelMessage(b, 1))
        assert op.servingMetadata == 2
        snap = op.snapshotState()
        assert sorted(snap.decode().splitlines()) == [f"{a}_1\t/models/a.pmml", f"{a}_2\t/models/a2.pmml"]
        # restore = union of every subtask's snapshot (getUnionListState)
        other = f"{b}_5\t/models/b5.pmml\n".encode()
        op2 = api.EvaluationCoFunction(None)
        op2.initializeState([snap, other, snap])
        assert op2.servingMetadata == 3 and op2.servingModels == 0            # models are not checkpointed
        op2.close()
    finally:
        op.close()


def test_host_symbols_are_bound():
    from flink_jpmml_b200 import capi
    import ctypes
    lib = ctypes.CDLL(capi.LIB_PATH)
    assert all(hasattr(lib, s) for s in api.HOST_EXPORTS)
