/*
B200 engine settings of a Model.  Embedded in ModelSpec as `b200` (patches/0001-modelspec-b200.patch); when absent the
operator behaves exactly as before and deploys the stock ollama/ollama server.
*/
package v1

// B200Spec selects the B200-native GGUF engine for the model's server pods and sizes its GPU request.
type B200Spec struct {
	// Number of B200 GPUs each replica asks the device plugin for (resources.limits["nvidia.com/gpu"]).
	// Replicas stay the data-parallel width (spec.replicas); this is the width of ONE replica. Defaults to 1.
	// +optional
	// +kubebuilder:validation:Minimum=1
	// +kubebuilder:validation:Maximum=8
	GPUs *int32 `json:"gpus,omitempty"`
	// Tensor-parallel degree inside a replica: the engine splits every layer over this many of the replica's GPUs
	// (NVLink peer memory, no collective library). Must divide the model's head counts; 1, 2, 4 or 8.
	// Defaults to the value of gpus.
	// +optional
	// +kubebuilder:validation:Enum=1;2;4;8
	TensorParallel *int32 `json:"tensorParallel,omitempty"`
	// Image of the engine's server (the process that answers the Ollama HTTP API from the read-only model store).
	// The puller init-container and the model store keep ollama/ollama.
	// +optional
	ServerImage string `json:"serverImage,omitempty"`
	// Upper bound of concurrently decoded sequences per replica (the engine's batch slots). Defaults to 8.
	// +optional
	// +kubebuilder:validation:Minimum=1
	// +kubebuilder:validation:Maximum=32
	MaxBatch *int32 `json:"maxBatch,omitempty"`
}

// DeepCopyInto copies the receiver into out (hand-written; `make generate` would emit the same).
func (in *B200Spec) DeepCopyInto(out *B200Spec) {
	*out = *in
	if in.GPUs != nil {
		v := *in.GPUs
		out.GPUs = &v
	}
	if in.TensorParallel != nil {
		v := *in.TensorParallel
		out.TensorParallel = &v
	}
	if in.MaxBatch != nil {
		v := *in.MaxBatch
		out.MaxBatch = &v
	}
}

// DeepCopy returns a deep copy of the receiver, or nil.
func (in *B200Spec) DeepCopy() *B200Spec {
	if in == nil {
		return nil
	}
	out := new(B200Spec)
	in.DeepCopyInto(out)
	return out
}

// GPUCount is the per-replica GPU request with its default.
func (in *B200Spec) GPUCount() int32 {
	if in == nil || in.GPUs == nil {
		return 1
	}
	return *in.GPUs
}

// TensorParallelDegree is the split inside a replica with its default (all of the replica's GPUs).
func (in *B200Spec) TensorParallelDegree() int32 {
	if in == nil {
		return 1
	}
	if in.TensorParallel == nil {
		return in.GPUCount()
	}
	return *in.TensorParallel
}
