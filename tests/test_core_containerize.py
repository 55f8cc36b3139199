"""Port of core/tests/unit/containerize_test.py: instead of Dockerfile lines and docker/Cloud Build
mocks, the staged job image is checked - Jobfile lines, the file-path map (same rules as upstream
:254-351), the manifest, both builders, and the requirements check that replaces `pip install`."""
import os
import tarfile

import pytest

from cloud_b200.core import containerize, machine_config

CFG = machine_config.COMMON_MACHINE_CONFIGS
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(autouse=True)
def _env(monkeypatch, tmp_path):
    monkeypatch.chdir(os.path.join(HERE, "testdata"))
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))


def _builder(cls=containerize.LocalContainerBuilder, **over):
    kw = dict(entry_point="mnist_example_using_fit.py", preprocessed_entry_point=None,
              chief_config=CFG["K80_1X"], worker_config=CFG["K80_1X"],
              docker_registry="local/my-project", project_id="my-project")
    kw.update(over)
    return cls(**kw)


def test_jobfile_defaults():
    b = _builder()
    lines = b._create_docker_file()
    assert lines[0].startswith("FROM cloud_b200:") and lines[0].rstrip().endswith("-gpu")
    assert lines[1:] == ["WORKDIR /app/\n", "COPY /app/ /app/\n",
                         'ENTRYPOINT ["python", "mnist_example_using_fit.py"]\n']
    os.remove(b.docker_file_path)


def test_jobfile_with_requirements_destination_dir_cpu_and_custom_image():
    b = _builder(requirements_txt="requirements.txt", destination_dir="/my_app/temp/",
                 chief_config=CFG["CPU"], preprocessed_entry_point="/tmp/wrapper_abc.py")
    lines = b._create_docker_file()
    assert not lines[0].rstrip().endswith("-gpu")
    assert lines[1:] == ["WORKDIR /my_app/temp/\n",
                         "COPY /my_app/temp/requirements.txt /my_app/temp/requirements.txt\n",
                         "RUN check-requirements /my_app/temp/requirements.txt\n",
                         "COPY /my_app/temp/ /my_app/temp/\n",
                         'ENTRYPOINT ["python", "wrapper_abc.py"]\n']
    os.remove(b.docker_file_path)
    b2 = _builder(docker_base_image="my/custom:1")
    assert b2._create_docker_file()[0] == "FROM my/custom:1\n"
    os.remove(b2.docker_file_path)


def test_jobfile_tpu_worker_note():
    b = _builder(chief_config=CFG["CPU"], worker_config=CFG["TPU"])
    assert any("TPU workers are not available" in l for l in b._create_docker_file())
    os.remove(b.docker_file_path)


def test_file_path_map_defaults_and_variants(tmp_path):
    b = _builder()
    b._create_docker_file()
    assert b._get_file_path_map() == {".": "/app/", b.docker_file_path: "Jobfile"}
    os.remove(b.docker_file_path)

    wrapper = str(tmp_path / "wrapper.py")
    open(wrapper, "w").write("print(1)\n")
    b = _builder(preprocessed_entry_point=wrapper, requirements_txt="requirements.txt",
                 destination_dir="/my_app/temp/")
    b._create_docker_file()
    assert b._get_file_path_map() == {".": "/my_app/temp/", wrapper: "/my_app/temp/wrapper.py",
                                      "requirements.txt": "/my_app/temp/requirements.txt",
                                      b.docker_file_path: "Jobfile"}
    os.remove(b.docker_file_path)

    b = _builder(entry_point="sub/dir/train.py")
    b._create_docker_file()
    assert b._get_file_path_map() == {"sub/dir": "/app/", b.docker_file_path: "Jobfile"}
    os.remove(b.docker_file_path)

    b = _builder(entry_point=None, preprocessed_entry_point=wrapper, called_from_notebook=True)
    b._create_docker_file()
    assert b._get_file_path_map() == {wrapper: "/app/wrapper.py", b.docker_file_path: "Jobfile"}
    os.remove(b.docker_file_path)


def test_tarball_dereferences_symlinks(tmp_path, monkeypatch):
    src = tmp_path / "proj"
    src.mkdir()
    (src / "train.py").write_text("print('hi')\n")
    (tmp_path / "shared.txt").write_text("payload")
    os.symlink(tmp_path / "shared.txt", src / "link.txt")
    monkeypatch.chdir(src)
    b = _builder(entry_point="train.py")
    b._get_tar_file_path()
    with tarfile.open(b.tar_file_path) as tar:
        member = tar.getmember("app/link.txt")
        assert member.isfile() and tar.extractfile(member).read() == b"payload"
        assert "Jobfile" in tar.getnames()
    for f in b.get_generated_files():
        os.remove(f)


def test_local_builder_stages_image_and_manifest():
    b = _builder(requirements_txt="requirements.txt")
    uri = b.get_docker_image()
    assert uri.startswith("local/my-project/tf_cloud_train:")
    manifest = containerize.resolve_image(uri)
    assert manifest["entrypoint"] == ["python", "mnist_example_using_fit.py"]
    assert os.path.isfile(os.path.join(manifest["workdir"], "mnist_example_using_fit.py"))
    assert os.path.isfile(os.path.join(manifest["workdir"], "requirements.txt"))
    assert os.path.isfile(os.path.join(os.path.dirname(manifest["workdir"]), "Jobfile"))
    assert len(b.get_generated_files()) == 2


def test_cloud_builder_goes_through_bucket(tmp_path):
    b = _builder(containerize.CloudContainerBuilder, docker_image_bucket_name="test_gcs_bucket")
    uri = b.get_docker_image(max_status_check_attempts=3, delay_between_status_checks=0)
    bucket = os.path.join(os.environ["TFC_JOB_ROOT"], "buckets", "test_gcs_bucket")
    assert len(os.listdir(bucket)) == 1 and os.listdir(bucket)[0].startswith("tf_cloud_train_tar_")
    req = b._create_cloud_build_request_dict(uri, os.listdir(bucket)[0])
    assert req["projectId"] == "my-project" and req["images"] == [[uri]]
    assert req["source"]["storageSource"]["bucket"] == "test_gcs_bucket"
    assert os.path.isdir(containerize.resolve_image(uri)["workdir"])


def test_cloud_builder_reports_build_failure(monkeypatch):
    b = _builder(containerize.CloudContainerBuilder, docker_image_bucket_name="bkt")

    def broken(*a, **k):
        raise tarfile.TarError("corrupt context")
    monkeypatch.setattr(tarfile, "open", broken)
    b.tar_file_path = __file__
    b._get_tar_file_path = lambda: None
    monkeypatch.setattr(b, "_upload_tar_to_gcs", lambda: "obj")
    with pytest.raises(RuntimeError, match=r"error executing the staging build. Job status: FAILURE"):
        b.get_docker_image(max_status_check_attempts=2, delay_between_status_checks=0)


def test_missing_requirement_fails_staging(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    (tmp_path / "train.py").write_text("print(1)\n")
    (tmp_path / "requirements.txt").write_text("numpy>=1.0\nthis-package-does-not-exist==9.9\n# comment\n")
    b = _builder(entry_point="train.py", requirements_txt="requirements.txt")
    with pytest.raises(RuntimeError, match=r"requirements not installed.*this-package-does-not-exist"):
        b.get_docker_image()
