"""Custom training loop under a user-managed strategy (reference:
core/tests/testdata/mnist_example_using_ctl.py): `strategy.scope()`, a distributed dataset, a
per-replica batch of 64, per-example losses averaged over the GLOBAL batch, `GradientTape` +
`optimizer.apply_gradients`, `strategy.run` / `strategy.reduce(SUM)`, and an evaluation loop with
stand-alone metric objects.  `apply_gradients` is where the replicas meet: the gradients are summed
by the fused all-reduce + Adam kernels."""
import numpy as np

import cloud_b200 as tfc

keras = tfc.keras

(train_images, train_labels), (test_images, test_labels) = keras.datasets.fashion_mnist.load_data()
train_images, train_labels = train_images[:4096, ..., None], train_labels[:4096]
test_images, test_labels = test_images[:1024, ..., None], test_labels[:1024]

# Getting images in [0, 1] range.
train_images = train_images / np.float32(255)
test_images = test_images / np.float32(255)

strategy = tfc.distribute.experimental.MultiWorkerMirroredStrategy()
print("Number of devices: {}".format(strategy.num_replicas_in_sync))

BUFFER_SIZE = len(train_images)
BATCH_SIZE_PER_REPLICA = 64
GLOBAL_BATCH_SIZE = BATCH_SIZE_PER_REPLICA * strategy.num_replicas_in_sync
EPOCHS = 2

train_dataset = tfc.data.Dataset.from_tensor_slices((train_images, train_labels)) \
    .shuffle(BUFFER_SIZE, seed=0).batch(GLOBAL_BATCH_SIZE)
test_dataset = tfc.data.Dataset.from_tensor_slices((test_images, test_labels)).batch(GLOBAL_BATCH_SIZE)

train_dist_dataset = strategy.experimental_distribute_dataset(train_dataset)
test_dist_dataset = strategy.experimental_distribute_dataset(test_dataset)


def create_model():
    return keras.Sequential([
        keras.layers.Conv2D(32, 3, activation="relu"),
        keras.layers.MaxPooling2D(),
        keras.layers.Conv2D(64, 3, activation="relu"),
        keras.layers.MaxPooling2D(),
        keras.layers.Flatten(),
        keras.layers.Dense(64, activation="relu"),
        keras.layers.Dense(10, activation="softmax"),
    ])


with strategy.scope():
    # Reduction `none`: the reduction happens afterwards, dividing by the global batch size.
    loss_object = keras.losses.SparseCategoricalCrossentropy(reduction=keras.losses.Reduction.NONE)

    def compute_loss(labels, predictions):
        per_example_loss = loss_object(labels, predictions)
        return tfc.nn.compute_average_loss(per_example_loss, global_batch_size=GLOBAL_BATCH_SIZE)

    test_loss = keras.metrics.Mean(name="test_loss")
    train_accuracy = keras.metrics.SparseCategoricalAccuracy(name="train_accuracy")
    test_accuracy = keras.metrics.SparseCategoricalAccuracy(name="test_accuracy")

    model = create_model()
    optimizer = keras.optimizers.Adam()


def train_step(inputs):
    images, labels = inputs
    with tfc.GradientTape() as tape:
        predictions = model(images, training=True)
        loss = compute_loss(labels, predictions)
    gradients = tape.gradient(loss, model.trainable_variables)
    optimizer.apply_gradients(zip(gradients, model.trainable_variables))
    train_accuracy.update_state(labels, predictions)
    return loss


def test_step(inputs):
    images, labels = inputs
    predictions = model(images, training=False)
    t_loss = loss_object(labels, predictions)
    test_loss.update_state(t_loss)
    test_accuracy.update_state(labels, predictions)


def distributed_train_step(dataset_inputs):
    per_replica_losses = strategy.run(train_step, args=(dataset_inputs,))
    return strategy.reduce(tfc.distribute.ReduceOp.SUM, per_replica_losses, axis=None)


def distributed_test_step(dataset_inputs):
    return strategy.run(test_step, args=(dataset_inputs,))


for epoch in range(EPOCHS):
    total_loss = 0.0
    num_batches = 0
    for batch in train_dist_dataset:
        total_loss += float(distributed_train_step(batch))
        num_batches += 1
    train_loss = total_loss / num_batches

    for batch in test_dist_dataset:
        distributed_test_step(batch)

    if strategy.is_chief:
        print("Epoch {}: loss {:.4f} accuracy {:.4f} test loss {:.4f} test accuracy {:.4f}".format(
            epoch + 1, train_loss, train_accuracy.result(), test_loss.result(), test_accuracy.result()))
    test_loss.reset_states()
    train_accuracy.reset_states()
    test_accuracy.reset_states()
