"""Local filesystem client (reference: python/paddle/distributed/fleet/utils/fs.py LocalFS; the HDFS client is out
of scope)."""
import os
import shutil


class LocalFS:
    def ls_dir(self, fs_path):
        if not self.is_exist(fs_path):
            return [], []
        dirs, files = [], []
        for f in os.listdir(fs_path):
            (dirs if os.path.isdir(os.path.join(fs_path, f)) else files).append(f)
        return dirs, files

    def mkdirs(self, fs_path):
        os.makedirs(fs_path, exist_ok=True)

    def rename(self, src, dst):
        os.rename(src, dst)

    def delete(self, fs_path):
        if os.path.isdir(fs_path):
            shutil.rmtree(fs_path)
        elif os.path.exists(fs_path):
            os.remove(fs_path)

    def is_file(self, fs_path):
        return os.path.isfile(fs_path)

    def is_dir(self, fs_path):
        return os.path.isdir(fs_path)

    def is_exist(self, fs_path):
        return os.path.exists(fs_path)

    def touch(self, fs_path, exist_ok=True):
        open(fs_path, "a").close()

    def mv(self, src, dst, overwrite=False, test_exists=False):
        if overwrite and os.path.exists(dst):
            self.delete(dst)
        shutil.move(src, dst)

    def list_dirs(self, fs_path):
        return self.ls_dir(fs_path)[0]


class ExecuteError(Exception):
    pass


class FSFileExistsError(Exception):
    pass


class FSFileNotExistsError(Exception):
    pass


class FSTimeOut(Exception):
    pass


class HDFSClient:
    """HDFS through the ``hadoop fs`` command line (reference: fleet/utils/fs.py HDFSClient).  ``hadoop_home`` must hold
    ``bin/hadoop``; ``configs`` are passed as ``-D key=value``.  Every call shells out, retried until ``time_out`` ms."""

    def __init__(self, hadoop_home, configs=None, time_out=5 * 60 * 1000, sleep_inter=1000):
        import os
        self._base_cmd = [os.path.join(hadoop_home, "bin", "hadoop"), "fs"]
        for k, v in (configs or {}).items():
            self._base_cmd += ["-D", f"{k}={v}"]
        self._time_out, self._sleep_inter = time_out, sleep_inter
        self._hadoop = self._base_cmd[0]

    def _run(self, args, retry=True):
        import os
        import subprocess
        import time
        if not os.path.exists(self._hadoop):
            raise ExecuteError(f"{self._hadoop} not found: HDFSClient needs a Hadoop installation")
        deadline = time.time() + self._time_out / 1000.0
        while True:
            proc = subprocess.run(self._base_cmd + args, capture_output=True, text=True)
            if proc.returncode == 0 or not retry:
                return proc.returncode, proc.stdout.splitlines()
            if time.time() >= deadline:
                raise FSTimeOut(f"hadoop fs {' '.join(args)} kept failing for {self._time_out} ms: {proc.stderr.strip()[-200:]}")
            time.sleep(self._sleep_inter / 1000.0)

    def is_exist(self, fs_path):
        return self._run(["-test", "-e", fs_path], retry=False)[0] == 0

    def is_dir(self, fs_path):
        return self._run(["-test", "-d", fs_path], retry=False)[0] == 0

    def is_file(self, fs_path):
        return self.is_exist(fs_path) and not self.is_dir(fs_path)

    def ls_dir(self, fs_path):
        """(sub-directory names, file names)."""
        if not self.is_exist(fs_path):
            return [], []
        _, lines = self._run(["-ls", fs_path])
        dirs, files = [], []
        for line in lines:
            parts = line.split()
            if len(parts) != 8:
                continue
            name = parts[7].rstrip("/").split("/")[-1]
            (dirs if parts[0].startswith("d") else files).append(name)
        return dirs, files

    def list_dirs(self, fs_path):
        return self.ls_dir(fs_path)[0]

    def mkdirs(self, fs_path):
        if not self.is_exist(fs_path):
            self._run(["-mkdir", "-p", fs_path])

    def delete(self, fs_path):
        if self.is_exist(fs_path):
            self._run(["-rm", "-r", "-f", fs_path])

    def upload(self, local_path, fs_path, multi_processes=1, overwrite=False):
        import os
        if not os.path.exists(local_path):
            raise FSFileNotExistsError(f"{local_path} not exists")
        if self.is_exist(fs_path):
            if not overwrite:
                raise FSFileExistsError(f"{fs_path} exists")
            self.delete(fs_path)
        self._run(["-put", local_path, fs_path])

    def download(self, fs_path, local_path, multi_processes=1, overwrite=False):
        import os
        import shutil
        if not self.is_exist(fs_path):
            raise FSFileNotExistsError(f"{fs_path} not exits")
        if os.path.exists(local_path):
            if not overwrite:
                raise FSFileExistsError(f"{local_path} exists")
            shutil.rmtree(local_path) if os.path.isdir(local_path) else os.remove(local_path)
        self._run(["-get", fs_path, local_path])

    def mv(self, fs_src_path, fs_dst_path, overwrite=False, test_exists=True):
        if overwrite and self.is_exist(fs_dst_path):
            self.delete(fs_dst_path)
        if test_exists:
            if not self.is_exist(fs_src_path):
                raise FSFileNotExistsError(f"{fs_src_path} is not exists")
            if self.is_exist(fs_dst_path):
                raise FSFileExistsError(f"{fs_dst_path} exists already")
        self._run(["-mv", fs_src_path, fs_dst_path])

    rename = mv

    def touch(self, fs_path, exist_ok=True):
        if self.is_exist(fs_path):
            if exist_ok:
                return
            raise FSFileExistsError(fs_path)
        self._run(["-touchz", fs_path])

    def cat(self, fs_path=None):
        if not self.is_file(fs_path):
            return ""
        return "\n".join(self._run(["-cat", fs_path])[1])

    def need_upload_download(self):
        return True


class DistributedInfer:
    """Parameter-server inference helper of the reference (pulls sparse tables from PS nodes).  The parameter-server
    runtime is out of this framework's scope (collective training only), so constructing it says so."""

    def __init__(self, main_program=None, startup_program=None):
        raise NotImplementedError("DistributedInfer belongs to the parameter-server mode, which this framework does not implement "
                                  "(collective training only); use paddle.inference for serving")
