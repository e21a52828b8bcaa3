package model

import (
	"strconv"

	corev1 "k8s.io/api/core/v1"
	"k8s.io/apimachinery/pkg/api/resource"

	ollamav1 "github.com/nekomeowww/ollama-operator/api/v1"
)

const (
	// B200ServerImage is the default image of the B200-native server (server/b2-server + libb200gguf.so).
	B200ServerImage = "ghcr.io/ollama-operator/b200-gguf-server:latest"
	// GPUResourceName is what the NVIDIA device plugin advertises.
	GPUResourceName corev1.ResourceName = "nvidia.com/gpu"
)

// NewB200ServerContainer is NewOllamaServerContainer for the B200 engine: the same contract towards kubelet and
// clients -- argv `serve`, OLLAMA_HOST, port "ollama" 11434, the image-storage volume at /root/.ollama, both probes
// on GET /api/tags -- with the engine's image, a GPU limit and the engine's sizing passed through the environment.
func NewB200ServerContainer(readOnly bool, spec *ollamav1.B200Spec) corev1.Container {
	c := NewOllamaServerContainer(readOnly)
	c.Image = B200ServerImage
	if spec != nil && spec.ServerImage != "" {
		c.Image = spec.ServerImage
	}
	gpus := spec.GPUCount()
	qty := resource.MustParse(strconv.Itoa(int(gpus)))
	c.Resources = corev1.ResourceRequirements{
		Limits:   corev1.ResourceList{GPUResourceName: qty},
		Requests: corev1.ResourceList{GPUResourceName: qty},
	}
	c.Env = append(c.Env,
		corev1.EnvVar{Name: "OLLAMA_MODELS", Value: "/root/.ollama/models"},
		corev1.EnvVar{Name: "B2_TP_SIZE", Value: strconv.Itoa(int(spec.TensorParallelDegree()))},
	)
	if spec != nil && spec.MaxBatch != nil {
		c.Env = append(c.Env, corev1.EnvVar{Name: "B2_MAX_BATCH", Value: strconv.Itoa(int(*spec.MaxBatch))})
	}
	// Loading a 70B model takes longer than Ollama's lazy first-request load; the probes answer from the
	// first second either way (GET /api/tags touches neither the GPU nor the engine thread).
	return c
}

// ServerContainerFor picks the server container of a Model: the B200 engine when spec.b200 is set, else the stock one.
func ServerContainerFor(spec ollamav1.ModelSpec, readOnly bool) corev1.Container {
	if spec.B200 != nil {
		return NewB200ServerContainer(readOnly, spec.B200)
	}
	return NewOllamaServerContainer(readOnly)
}
