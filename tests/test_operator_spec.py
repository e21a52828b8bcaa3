"""Operator-side changes (operator/) cannot be compiled here (no Go toolchain), so this checks what can be
checked from Python: the golden server container keeps the reference's container contract (pkg/model/pod.go:14-66:
argv, env, port, mount, both probes on /api/tags) and adds the GPU request; the CRD schema addition matches the JSON tags
of the Go type; the Go builder sets the fields the golden shows; and the server binary implements that contract."""
import os
import re

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OP = os.path.join(ROOT, "operator")


def _read(*p):
    return open(os.path.join(OP, *p)).read()


def test_golden_container_keeps_the_reference_contract_and_adds_gpus():
    c = yaml.safe_load(_read("testdata", "server_container_golden.yaml"))
    assert c["name"] == "server" and c["args"] == ["serve"]
    env = {e["name"]: e["value"] for e in c["env"]}
    assert env["OLLAMA_HOST"] == "0.0.0.0" and env["OLLAMA_MODELS"] == "/root/.ollama/models" and env["B2_TP_SIZE"] == "2"
    assert c["ports"] == [{"name": "ollama", "protocol": "TCP", "containerPort": 11434}]
    assert c["volumeMounts"] == [{"name": "image-storage", "mountPath": "/root/.ollama", "readOnly": True}]
    for probe, timeout in (("readinessProbe", 5), ("livenessProbe", 1)):
        p = c[probe]
        assert p["httpGet"] == {"path": "/api/tags", "port": "ollama"} and p["timeoutSeconds"] == timeout
        assert p["initialDelaySeconds"] == 15 and p["failureThreshold"] == 2500 and p["successThreshold"] == 1
    assert c["resources"]["limits"] == {"nvidia.com/gpu": "2"} == c["resources"]["requests"]
    assert c["image"] != "ollama/ollama"


def test_crd_properties_match_the_go_json_tags():
    crd = yaml.safe_load(_read("config", "crd", "b200_properties.yaml"))["b200"]
    go = _read("api", "v1", "model_types_b200.go")
    tags = re.findall(r'`json:"(\w+),omitempty"`', go)
    assert sorted(tags) == sorted(crd["properties"]) == ["gpus", "maxBatch", "serverImage", "tensorParallel"]
    assert crd["properties"]["gpus"]["minimum"] == 1 and crd["properties"]["gpus"]["maximum"] == 8
    assert crd["properties"]["tensorParallel"]["enum"] == [1, 2, 4, 8]
    for f in ("GPUs", "TensorParallel", "MaxBatch"):            # every pointer field is deep-copied
        assert f"in.{f} != nil" in go
    patch = _read("patches", "0001-modelspec-b200.patch")
    assert 'B200 *B200Spec `json:"b200,omitempty"`' in patch and "in.B200 != nil" in patch


def test_go_builder_sets_what_the_golden_shows_and_the_server_reads_it():
    pod = _read("pkg", "model", "pod_b200.go")
    golden = yaml.safe_load(_read("testdata", "server_container_golden.yaml"))
    assert f'B200ServerImage = "{golden["image"]}"' in pod
    assert 'GPUResourceName corev1.ResourceName = "nvidia.com/gpu"' in pod
    for name in ("OLLAMA_MODELS", "B2_TP_SIZE", "B2_MAX_BATCH"):
        assert f'Name: "{name}"' in pod
    assert "NewOllamaServerContainer(readOnly)" in pod           # the reference's container is the base: contract inherited, not restated
    server = open(os.path.join(ROOT, "server", "b2_server.cpp")).read()
    for name in ("OLLAMA_HOST", "OLLAMA_MODELS", "B2_TP_SIZE", "B2_MAX_BATCH"):
        assert f'getenv("{name}")' in server
    assert '"/api/tags"' in server and "11434" in server and 'strcmp(argv[1], "serve")' in server
    sync = _read("pkg", "model", "deployment_b200.go")
    assert "containers[i].Image != want.Image" in sync and "containers[i].Resources" in sync
