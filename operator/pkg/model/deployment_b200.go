package model

import (
	appsv1 "k8s.io/api/apps/v1"
	corev1 "k8s.io/api/core/v1"
	"k8s.io/apimachinery/pkg/api/equality"

	ollamav1 "github.com/nekomeowww/ollama-operator/api/v1"
)

// SyncServerContainer brings the "server" container of an existing Deployment in line with the Model's spec and
// reports whether anything changed.  UpdateDeployment reconciles replicas only; without this an edit of spec.b200
// (another image, more GPUs per replica) would never reach the pods.
func SyncServerContainer(deployment *appsv1.Deployment, spec ollamav1.ModelSpec) bool {
	want := ServerContainerFor(spec, true)
	containers := deployment.Spec.Template.Spec.Containers
	for i := range containers {
		if containers[i].Name != want.Name {
			continue
		}
		changed := false
		if containers[i].Image != want.Image {
			containers[i].Image = want.Image
			changed = true
		}
		if !equality.Semantic.DeepEqual(containers[i].Resources, want.Resources) {
			containers[i].Resources = want.Resources
			changed = true
		}
		if !equality.Semantic.DeepEqual(containers[i].Env, want.Env) {
			containers[i].Env = append([]corev1.EnvVar(nil), want.Env...)
			changed = true
		}
		return changed
	}
	return false
}
